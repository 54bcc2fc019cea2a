N=8
export PYTHONPATH=.
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus $N --steps 30 --warmup 3 --no-cpu > gpurun_out/r2_bench_n8.json 2> gpurun_out/r2_bench_n8.err
echo "bench n$N rc=$? stdout lines=$(wc -l < gpurun_out/r2_bench_n8.json)"; cut -c1-400 gpurun_out/r2_bench_n8.json
