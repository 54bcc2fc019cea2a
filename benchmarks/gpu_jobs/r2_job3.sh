set -x
export PYTHONPATH=.
( timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_job3_gputests.log 2>&1
echo "gputests rc=$?"
tail -5 gpurun_out/r2_job3_gputests.log
cd benchmarks/micro
for cells in 64 256 1024 4096; do for skew in 0 1; do ./smem_incr $cells 256 8 $skew sweep; ./smem_incr $cells 32 32 $skew sweep; done; done > ../../gpurun_out/r2_micro_replication.txt 2>&1
cd ../..
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r2_job3_bench.json 2> gpurun_out/r2_job3_bench.err
echo "bench rc=$?"; tail -3 gpurun_out/r2_job3_bench.err
timeout 300 python benchmarks/make_frontier.py c5 c4 c2 --out gpurun_out/frontiers
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_job3_bench_ref.json 2> gpurun_out/r2_job3_bench_ref.err
echo "ref rc=$?"; tail -3 gpurun_out/r2_job3_bench_ref.err
