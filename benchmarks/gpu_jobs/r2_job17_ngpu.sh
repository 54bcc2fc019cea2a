# usage: bash benchmarks/gpu_jobs/r2_job17_ngpu.sh <N>   (under gpurun --gpus N)
N=${1:-2}
export PYTHONPATH=.
nvidia-smi -L | head -8
if [ "$N" = "2" ]; then
  ( timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q ) > gpurun_out/r2_n${N}_multitest.log 2>&1
  echo "multitest rc=$?"; tail -6 gpurun_out/r2_n${N}_multitest.log
fi
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 900 $TR bench.py --gpus $N --steps 30 --warmup 3 --no-cpu > gpurun_out/r2_bench_n${N}.json 2> gpurun_out/r2_bench_n${N}.err
echo "bench n$N rc=$?"; tail -3 gpurun_out/r2_bench_n${N}.err; cut -c1-2500 gpurun_out/r2_bench_n${N}.json
timeout 600 $TR benchmarks/run_configs.py c3 c4 c5 > gpurun_out/r2_configs_n${N}.jsonl 2> gpurun_out/r2_configs_n${N}.err
echo "configs rc=$?"
python - <<PY
import json
for l in open('gpurun_out/r2_configs_n${N}.jsonl'):
    if not l.startswith('{'): continue
    d=json.loads(l)
    print({k: d[k] for k in d if k in ('config','n_gpus','learn_ms','kernel_ms','ci_tests','tables','collectives','adj_sha','score_ms','dp_ms','gather_ms','learn_time_s','best_score')})
PY
