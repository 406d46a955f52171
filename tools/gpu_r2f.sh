cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
(time timeout 3000 python -m pytest tests -m gpu -q 2>&1 | tail -30) > gpurun_out/r2f_pytest.log 2>&1
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/r2f_bench_default.json 2> gpurun_out/r2f_bench_default.err
for wl in rectangular_7x7_1-32-1 sycamore_53_10 sycamore_53_16; do
  timeout 900 python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2f_bench_${wl}.json 2> gpurun_out/r2f_bench_${wl}.err
done
tail -12 gpurun_out/r2f_pytest.log
