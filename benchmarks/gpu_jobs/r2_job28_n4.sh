N=4
export PYTHONPATH=.
( timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_n4_gputests.log 2>&1
echo "gputests (incl. 2-GPU test) rc=$?"; tail -4 gpurun_out/r2_n4_gputests.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus $N --steps 30 --warmup 3 --no-cpu > gpurun_out/r2_bench_n4.json 2> gpurun_out/r2_bench_n4.err
echo "bench n$N rc=$? stdout lines=$(wc -l < gpurun_out/r2_bench_n4.json)"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench_n4.json'))
print('value',d['value'],'ms',d['ms_per_step'],'parity',d.get('parity_ok'))
print({k:round(v,2) for k,v in d['limiter'].items() if isinstance(v,(int,float))})
for s in d['secondary']: print({k:s[k] for k in s if k in ('workload','ms_per_step','learn_time_s','parity_ok')})
PY
