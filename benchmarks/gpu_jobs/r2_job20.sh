export PYTHONPATH=.
( timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_job20_gputests.log 2>&1
echo "gputests rc=$?"; tail -12 gpurun_out/r2_job20_gputests.log
O=gpurun_out/r2_job20_configs.jsonl; rm -f $O
run() { timeout 300 python benchmarks/run_configs.py "$@" >> $O 2>> gpurun_out/r2_job20_configs.err; }
run c2 c4 c5
run c2 c4 c5 opt:min_class=1
run c2 c4 c5 opt:min_class=2
run c2 c4 c5 opt:streams=1
run c2ns
python - <<'PY'
import json
for l in open('gpurun_out/r2_job20_configs.jsonl'):
    if not l.startswith('{'): continue
    d=json.loads(l)
    print(d['config'], d['options'], 'learn_ms=%.2f kernel_ms=%.2f tests=%d tables=%d launches=%d enum=%.2f eval=%.2f res=%.2f prep=%.2f exposed=%.2f sha=%d'%(d['learn_ms'],d['kernel_ms'],d['ci_tests'],d['tables'],d['launches'],d['enumerate_ms'],d['evaluate_ms'],d['resolve_ms'],d['host_prep_ms'],d['exposed_ms'] or 0,d['adj_sha']))
PY
