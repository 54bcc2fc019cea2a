set -x
export PYTHONPATH=.
( timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_job2_gputests.log 2>&1
echo "gputests rc=$?"
tail -5 gpurun_out/r2_job2_gputests.log
for tb in 8192 32768 1000000; do
  for ov in 0 1; do
    timeout 300 python benchmarks/run_configs.py c2 c4 c5 tb=$tb overlap=$ov >> gpurun_out/r2_job2_configs.jsonl 2>> gpurun_out/r2_job2_configs.err
  done
done
python - <<'PY'
import json
for l in open('gpurun_out/r2_job2_configs.jsonl'):
    if not l.startswith('{'): continue
    d=json.loads(l)
    print(d['config'], d['pc_kwargs'], 'learn_ms=%.2f kernel_ms=%.2f tests=%d tables=%d launches=%d enum=%.2f eval=%.2f res=%.2f prep=%.2f atoms_ceiling=%.2f sha=%d'%(d['learn_ms'],d['kernel_ms'],d['ci_tests'],d['tables'],d['launches'],d['enumerate_ms'],d['evaluate_ms'],d['resolve_ms'],d['host_prep_ms'],d['atoms_ceiling_ms'],d['adj_sha']), d['tests_per_depth'], d['batches_per_depth'])
PY
