export PYTHONPATH=.
timeout 900 python bench.py > gpurun_out/r2_bench_n1_b.json 2> gpurun_out/r2_bench_n1_b.err
echo "bench rc=$? lines=$(wc -l < gpurun_out/r2_bench_n1_b.json)"
timeout 300 python benchmarks/run_configs.py c3mp23 c1 > gpurun_out/r2_job26_es.jsonl 2>&1
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench_n1_b.json'))
print(d['value'], d['ms_per_step'])
for s in d['secondary']: print({k:s[k] for k in s if k in ('workload','learn_time_s','phases_ms','kernel_s','ms_per_step')})
for l in open('gpurun_out/r2_job26_es.jsonl'):
    if l.startswith('{'):
        e=json.loads(l); print({k:e[k] for k in e if k in ('config','learn_ms','kernel_ms','score_ms','dp_ms','max_parents')})
PY
