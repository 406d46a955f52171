set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv
(time timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -30) > gpurun_out/r2a_pytest.log 2>&1
for wl in sycamore_53_8_batch64 bristlecone_70_1-24-1 bristlecone_70_1-32-1 rectangular_7x7_1-32-1; do
  for fb in 0 62; do
    QFLEX_FUSE_BITS=$fb timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2a_bench_${wl}_fuse${fb}.json 2> gpurun_out/r2a_bench_${wl}_fuse${fb}.err
  done
done
tail -5 gpurun_out/r2a_pytest.log
