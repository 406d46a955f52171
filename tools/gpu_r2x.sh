cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
# the no-allocation entry point: alone in a fresh process, then after the whole suite
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r2x_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -15 gpurun_out/r2x_noalloc_alone.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2x_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r2x_smoke.log
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/r2x_gpu_suite.log 2>&1; echo "suite rc=$?"; tail -8 gpurun_out/r2x_gpu_suite.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2x_bench.json 2> gpurun_out/r2x_bench.err; echo "bench rc=$?"; cat gpurun_out/r2x_bench.json | cut -c1-600
