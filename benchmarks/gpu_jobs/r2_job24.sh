export PYTHONPATH=.
O=gpurun_out/r2_ab_streams.jsonl; rm -f $O
run() { timeout 300 python benchmarks/run_configs.py c2 c4 c5 "$@" >> $O 2>> gpurun_out/r2_ab_streams.err; }
run
run opt:streams=8
run opt:largest_first=1
run opt:streams=8 opt:largest_first=1
run opt:streams=6 opt:largest_first=1
python - <<'PY'
import json
for l in open('gpurun_out/r2_ab_streams.jsonl'):
    if not l.startswith('{'): continue
    d=json.loads(l)
    print(d['config'], d['options'], 'learn_ms=%.2f kernel_ms=%.2f launches=%d exposed=%.2f sha=%d'%(d['learn_ms'],d['kernel_ms'],d['launches'],d['exposed_ms'] or 0,d['adj_sha']))
PY
