cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
python tools/tc_bias.py > gpurun_out/r2c_tc_bias.log 2>&1
(timeout 1200 python -m pytest tests -m gpu -q -x tests/test_gpu_kernels.py tests/test_reference_frontends.py 2>&1 | tail -15) > gpurun_out/r2c_pytest.log 2>&1
for wl in sycamore_53_8_batch64 bristlecone_70_1-24-1 bristlecone_48_1-24-1; do
  timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2c_bench_${wl}.json 2> gpurun_out/r2c_bench_${wl}.err
done
cat gpurun_out/r2c_tc_bias.log; tail -5 gpurun_out/r2c_pytest.log
