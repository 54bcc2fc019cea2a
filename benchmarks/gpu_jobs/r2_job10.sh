export PYTHONPATH=.
timeout 300 python benchmarks/run_configs.py c5 reps=1 overlap=0 opt:debug=2 > gpurun_out/r2_job10_log.jsonl 2> gpurun_out/r2_job10_log.err
grep "obn batch" gpurun_out/r2_job10_log.err | tail -12 | cut -c1-700
timeout 300 python benchmarks/run_configs.py c5 overlap=0 opt:streams=1 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('streams=1', d['learn_ms'], d['kernel_ms'])"
