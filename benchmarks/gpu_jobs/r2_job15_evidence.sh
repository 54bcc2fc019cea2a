# Evidence pass with the final kernels: ncu --set full of one C5 / C4 pc() call, the launch list of a short bench run,
# then the default bench line and the reference arm (never under ncu).
export PYTHONPATH=.
for cfg in c5 c4; do
  timeout 900 ncu --set full --clock-control none -k "regex:count_kernel|ci_epilogue_kernel" -c 400 -o gpurun_out/r2_ncu_${cfg}_full \
      python benchmarks/run_configs.py $cfg reps=1 > gpurun_out/r2_ncu_${cfg}_full.out 2>&1
  echo "ncu $cfg rc=$?"
  ncu -i gpurun_out/r2_ncu_${cfg}_full.ncu-rep --page raw --csv > gpurun_out/r2_ncu_${cfg}_raw.csv 2>/dev/null
  rm -f gpurun_out/r2_ncu_${cfg}_full.ncu-rep
  python benchmarks/ncu_summary.py gpurun_out/r2_ncu_${cfg}_raw.csv $cfg 2 gpurun_out/r2_ncu_${cfg} > gpurun_out/r2_ncu_${cfg}_summary.txt 2>&1
  tail -12 gpurun_out/r2_ncu_${cfg}_summary.txt
done
cp profiles/r2_traffic.json gpurun_out/r2_traffic.json
# the bench itself (dram_frac now comes from the capture above)
timeout 1200 python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err
echo "bench rc=$?"; cut -c1-1500 gpurun_out/r2_bench_n1.json
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_reference.json 2> gpurun_out/r2_bench_reference.err
echo "reference rc=$?"; cut -c1-1200 gpurun_out/r2_bench_reference.json
# launch list of a short bench run
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r2_launches_bench.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-secondary --no-e2e > gpurun_out/r2_launches_bench.out 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/r2_launches_bench.csv
