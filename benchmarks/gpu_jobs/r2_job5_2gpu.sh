set -x
export PYTHONPATH=.
nvidia-smi -L
( timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q ) > gpurun_out/r2_job5_multitest.log 2>&1
echo "multitest rc=$?"; tail -15 gpurun_out/r2_job5_multitest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2_job5_bench_n2.json 2> gpurun_out/r2_job5_bench_n2.err
echo "bench n2 rc=$?"; tail -5 gpurun_out/r2_job5_bench_n2.err; cut -c1-1500 gpurun_out/r2_job5_bench_n2.json
timeout 300 $TR benchmarks/run_configs.py c4 c5 > gpurun_out/r2_job5_configs_n2.jsonl 2> gpurun_out/r2_job5_configs_n2.err
timeout 300 $TR benchmarks/run_configs.py c4 c5 hostcomm > gpurun_out/r2_job5_configs_n2_hostcomm.jsonl 2>> gpurun_out/r2_job5_configs_n2.err
timeout 300 $TR benchmarks/run_configs.py c4 c5 shard=0 >> gpurun_out/r2_job5_configs_n2.jsonl 2>> gpurun_out/r2_job5_configs_n2.err
timeout 300 $TR benchmarks/run_configs.py c4 c5 overlap=0 >> gpurun_out/r2_job5_configs_n2.jsonl 2>> gpurun_out/r2_job5_configs_n2.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_job5_configs_n2*.jsonl')):
  for l in open(f):
    if not l.startswith('{'): continue
    d=json.loads(l)
    print(f[-22:], d['config'], d.get('n_gpus'), d['pc_kwargs'], 'learn_ms=%.2f kernel_ms=%.2f tests=%d tables=%d enum=%.2f eval=%.2f res=%.2f prep=%.2f coll=%s sha=%d'%(d['learn_ms'],d['kernel_ms'],d['ci_tests'],d['tables'],d['enumerate_ms'],d['evaluate_ms'],d['resolve_ms'],d['host_prep_ms'],d['collectives'],d['adj_sha']))
PY
