export PYTHONPATH=.
( timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_final_gputests.log 2>&1
echo "gputests rc=$?"; tail -5 gpurun_out/r2_final_gputests.log
python __graft_entry__.py --smoke 2>&1 | tail -1
timeout 900 python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err
echo "bench rc=$? lines=$(wc -l < gpurun_out/r2_bench_n1.json)"; cut -c1-200 gpurun_out/r2_bench_n1.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_bench_reference.json 2> gpurun_out/r2_bench_reference.err
echo "reference rc=$? lines=$(wc -l < gpurun_out/r2_bench_reference.json)"; cut -c1-120 gpurun_out/r2_bench_reference.json
