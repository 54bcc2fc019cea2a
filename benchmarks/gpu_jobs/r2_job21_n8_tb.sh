N=8
export PYTHONPATH=.
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
O=gpurun_out/r2_ab_n8_target_batch.jsonl; rm -f $O
for tb in 16384 32768 65536; do
  timeout 300 $TR benchmarks/run_configs.py c4 c5 tb=$tb >> $O 2>> gpurun_out/r2_ab_n8_target_batch.err
done
python - <<'PY'
import json
for l in open('gpurun_out/r2_ab_n8_target_batch.jsonl'):
    if not l.startswith('{'): continue
    d=json.loads(l)
    print(d['config'], d['pc_kwargs'], 'learn_ms=%.2f kernel_ms=%.2f tests=%d tables=%d launches=%d enum=%.2f eval=%.2f res=%.2f exposed=%.2f coll=%s sha=%d'%(d['learn_ms'],d['kernel_ms'],d['ci_tests'],d['tables'],d['launches'],d['enumerate_ms'],d['evaluate_ms'],d['resolve_ms'],d['exposed_ms'] or 0,d['collectives'],d['adj_sha']))
PY
