cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
(timeout 2400 python -m pytest tests -m gpu -q -x 2>&1 | tail -15) > gpurun_out/r2k_pytest.log 2>&1
timeout 900 python bench.py --workload sycamore_53_16 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2k_bench_m16.json 2> gpurun_out/r2k_bench_m16.err
QFLEX_LINE_PROGRAMS=0 timeout 900 python bench.py --workload sycamore_53_16 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2k_bench_m16_noprog.json 2> gpurun_out/r2k_bench_m16_noprog.err
timeout 900 python bench.py --workload bristlecone_48_1-24-1 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2k_bench_b48.json 2> gpurun_out/r2k_bench_b48.err
timeout 1200 python bench.py --workload sycamore_53_14 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r2k_bench_m14.json 2> gpurun_out/r2k_bench_m14.err
tail -15 gpurun_out/r2k_pytest.log
