set -x
export PYTHONPATH=.
timeout 600 ncu --metrics gpu__time_duration.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared_op_atom.sum,smsp__inst_executed_op_shared_atom.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed,lts__t_sectors_op_read.sum,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:count_kernel -c 200 --csv --page raw --log-file gpurun_out/r2_job8_c5_launches.csv python benchmarks/run_configs.py c5 reps=1 > gpurun_out/r2_job8_c5.out 2>&1
tail -2 gpurun_out/r2_job8_c5.out
