cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2zr_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r2zr_smoke.log
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/r2zr_gpu_suite.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r2zr_gpu_suite.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2zr_bench.json 2> gpurun_out/r2zr_bench.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/r2zr_bench.json
