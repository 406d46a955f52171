cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
python tools/gpu_r2b.py > gpurun_out/r2b_precision.log 2>&1
(time timeout 3000 python -m pytest tests -m gpu -q 2>&1 | tail -40) > gpurun_out/r2b_pytest.log 2>&1
for wl in bristlecone_70_1-32-1 sycamore_53_8_batch64 bristlecone_70_1-24-1 bristlecone_48_1-24-1; do
  timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 > gpurun_out/r2b_bench_${wl}.json 2> gpurun_out/r2b_bench_${wl}.err
done
cat gpurun_out/r2b_precision.log; tail -15 gpurun_out/r2b_pytest.log
