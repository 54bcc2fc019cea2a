export PYTHONPATH=.
mkdir -p profiles
[ -f gpurun_out/r2_traffic.json ] && cp gpurun_out/r2_traffic.json profiles/r2_traffic.json
timeout 300 python benchmarks/run_configs.py c2 c4 c5 > gpurun_out/r2_configs_1gpu.jsonl 2> gpurun_out/r2_configs_1gpu.err
python - <<'PY'
import json
for l in open('gpurun_out/r2_configs_1gpu.jsonl'):
    if not l.startswith('{'): continue
    d=json.loads(l)
    print(d['config'], 'learn_ms=%.2f kernel_ms=%.2f tests=%d tables=%d launches=%d enum=%.2f eval=%.2f res=%.2f prep=%.2f exposed=%.2f wait=%.2f sha=%d'%(d['learn_ms'],d['kernel_ms'],d['ci_tests'],d['tables'],d['launches'],d['enumerate_ms'],d['evaluate_ms'],d['resolve_ms'],d['host_prep_ms'],d['exposed_ms'],d['collect_wait_ms'],d['adj_sha']))
PY
timeout 1200 python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err
echo "bench rc=$?"; tail -3 gpurun_out/r2_bench_n1.err; cut -c1-2500 gpurun_out/r2_bench_n1.json
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_reference.json 2> gpurun_out/r2_bench_reference.err
echo "reference rc=$?"; cut -c1-1200 gpurun_out/r2_bench_reference.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r2_launches_bench.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-secondary --no-e2e > gpurun_out/r2_launches_bench.out 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/r2_launches_bench.csv
