set -x
export PYTHONPATH=.
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
( OBN_TEST_PC_OVERLAP=1 OBN_TEST_CI_POOL_WIDE=1 OBN_TEST_CI_ZERO_COPY=1 timeout 400 python -m pytest tests/test_zz_gpu_ci_pool.py -m gpu -x -q ) > gpurun_out/r2_optin_tests.log 2>&1
echo "optin rc=$?"
timeout 300 python benchmarks/run_configs.py c2 c4 c5 > gpurun_out/r2_base_configs.jsonl 2> gpurun_out/r2_base_configs.err
OBN_PC_OVERLAP=1 timeout 300 python benchmarks/run_configs.py c4 c5 > gpurun_out/r2_overlap.jsonl 2>&1
OBN_CI_ZERO_COPY=1 timeout 300 python benchmarks/run_configs.py c2 c4 c5 > gpurun_out/r2_zc.jsonl 2>&1
OBN_CI_POOL=1 timeout 300 python benchmarks/run_configs.py c4 c5 > gpurun_out/r2_pool.jsonl 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r2_c5_launches.csv python benchmarks/run_configs.py c5 > gpurun_out/r2_c5_launches.out 2>&1
timeout 900 ncu --set full --clock-control none -k regex:count_kernel -c 75 -o gpurun_out/r2_c5_full python benchmarks/run_configs.py c5 > gpurun_out/r2_c5_full.out 2>&1
ncu -i gpurun_out/r2_c5_full.ncu-rep --page raw --csv > gpurun_out/r2_c5_full_raw.csv 2>/dev/null
ls -la gpurun_out
sz=$(stat -c %s gpurun_out/r2_c5_full.ncu-rep 2>/dev/null || echo 0)
if [ "$sz" -gt 40000000 ]; then rm -f gpurun_out/r2_c5_full.ncu-rep; fi
tail -3 gpurun_out/r2_optin_tests.log
cat gpurun_out/r2_base_configs.jsonl | cut -c1-600
