cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/r2zp_gpu_suite.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r2zp_gpu_suite.log
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2zp_bench.json 2> gpurun_out/r2zp_bench.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/r2zp_bench.json
QFLEX_PDL=0 timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2zp_bench_nopdl.json 2> gpurun_out/r2zp_bench_nopdl.err; echo "bench(no pdl) rc=$?"; cut -c1-200 gpurun_out/r2zp_bench_nopdl.json
for e in 1 0; do QFLEX_PDL=$e timeout 300 python bench.py --workload sycamore_53_8_batch64 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2zp_m8_pdl$e.json 2>/dev/null; echo "m8 pdl=$e"; cut -c1-120 gpurun_out/r2zp_m8_pdl$e.json; done
