export PYTHONPATH=.
( timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q ) > gpurun_out/r2_job14_static_alone.log 2>&1
echo "static alone rc=$?"
tail -15 gpurun_out/r2_job14_static_alone.log
( timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_job14_gputests.log 2>&1
echo "gputests rc=$?"
tail -8 gpurun_out/r2_job14_gputests.log
timeout 300 python benchmarks/run_configs.py c2 c4 c5 > gpurun_out/r2_job14_configs.jsonl 2> gpurun_out/r2_job14_configs.err
python - <<'PY'
import json
for l in open('gpurun_out/r2_job14_configs.jsonl'):
    if not l.startswith('{'): continue
    d=json.loads(l)
    print(d['config'], 'learn_ms=%.2f kernel_ms=%.2f tests=%d tables=%d launches=%d enum=%.2f eval=%.2f res=%.2f prep=%.2f sha=%d'%(d['learn_ms'],d['kernel_ms'],d['ci_tests'],d['tables'],d['launches'],d['enumerate_ms'],d['evaluate_ms'],d['resolve_ms'],d['host_prep_ms'],d['adj_sha']))
PY
