export PYTHONPATH=.
timeout 300 python benchmarks/run_configs.py c1 c2 c3 c4 c5 > gpurun_out/r2_configs_1gpu.jsonl 2> gpurun_out/r2_configs_1gpu.err
python - <<'PY'
import json
for l in open('gpurun_out/r2_configs_1gpu.jsonl'):
    if not l.startswith('{'): continue
    d=json.loads(l)
    if 'enumerate_ms' in d and 'exposed_ms' in d:
        print(d['config'], 'learn_ms=%.2f kernel_ms=%.2f tests=%d tables=%d launches=%d enum=%.2f eval=%.2f res=%.2f prep=%.2f exposed=%.2f wait=%.2f sha=%d'%(d['learn_ms'],d['kernel_ms'],d['ci_tests'],d['tables'],d['launches'],d['enumerate_ms'],d['evaluate_ms'],d['resolve_ms'],d['host_prep_ms'],d['exposed_ms'],d['collect_wait_ms'],d['adj_sha']))
    else:
        print({k: d[k] for k in d if k in ('config','learn_ms','kernel_ms','score_ms','dp_ms','gather_ms','local_scores')})
PY
