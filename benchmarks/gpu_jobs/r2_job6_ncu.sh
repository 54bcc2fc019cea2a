set -x
export PYTHONPATH=.
for cfg in c5 c4; do
timeout 600 ncu --set full --clock-control none -k regex:count_kernel -c 120 -o gpurun_out/r2_ncu_${cfg}_full python benchmarks/run_configs.py $cfg reps=1 > gpurun_out/r2_ncu_${cfg}_full.out 2>&1
ncu -i gpurun_out/r2_ncu_${cfg}_full.ncu-rep --page raw --csv > gpurun_out/r2_ncu_${cfg}_raw.csv 2>/dev/null
rm -f gpurun_out/r2_ncu_${cfg}_full.ncu-rep
done
ls -la gpurun_out | grep ncu
