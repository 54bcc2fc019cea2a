export PYTHONPATH=.
( timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_job22_gputests.log 2>&1
echo "gputests rc=$?"; tail -6 gpurun_out/r2_job22_gputests.log
python __graft_entry__.py --smoke 2>&1 | tail -2
timeout 300 python benchmarks/run_configs.py c1 c3mp23 > gpurun_out/r2_job22_configs.jsonl 2> gpurun_out/r2_job22_configs.err
python - <<'PY'
import json
for l in open('gpurun_out/r2_job22_configs.jsonl'):
    if not l.startswith('{'): continue
    d=json.loads(l)
    print({k: d[k] for k in d if k in ('config','learn_ms','kernel_ms','score_ms','dp_ms','gather_ms','local_scores','max_parents')})
PY
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/r2_job22_bench.json 2> gpurun_out/r2_job22_bench.err
echo "bench rc=$? lines=$(wc -l < gpurun_out/r2_job22_bench.json)"; cut -c1-300 gpurun_out/r2_job22_bench.json
