export PYTHONPATH=.
( timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_job12_gputests.log 2>&1
echo "gputests rc=$?"
tail -5 gpurun_out/r2_job12_gputests.log
O=gpurun_out/r2_job12_configs.jsonl
run() { timeout 300 python benchmarks/run_configs.py c2 c4 c5 "$@" >> $O 2>> gpurun_out/r2_job12_configs.err; }
run
run opt:pack4=0
timeout 300 python benchmarks/run_configs.py c5 reps=1 overlap=0 opt:debug=2 > /dev/null 2> gpurun_out/r2_job12_log.err
grep "obn batch" gpurun_out/r2_job12_log.err | tail -10 | cut -c1-600
python - <<'PY'
import json
for l in open('gpurun_out/r2_job12_configs.jsonl'):
    if not l.startswith('{'): continue
    d=json.loads(l)
    print(d['config'], d['pc_kwargs'], d['options'], 'learn_ms=%.2f kernel_ms=%.2f tests=%d tables=%d launches=%d enum=%.2f eval=%.2f res=%.2f prep=%.2f sha=%d'%(d['learn_ms'],d['kernel_ms'],d['ci_tests'],d['tables'],d['launches'],d['enumerate_ms'],d['evaluate_ms'],d['resolve_ms'],d['host_prep_ms'],d['adj_sha']))
PY
