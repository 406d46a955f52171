cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
export QFLEX_KB_NOPERM=1
timeout 600 python -m pytest tests/test_gpu_tensorcore.py tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -2
{
echo "dbg=0"; timeout 120 python tools/kernel_bench.py 24,4,4 2>&1 | grep -E '"gemm"'
for d in 256 335 271; do echo "dbg=$d 24,4,4"; QFLEX_FUSED_DBG=$d timeout 120 python tools/kernel_bench.py 24,4,4 2>&1 | grep -E '"gemm"|pair-prof' | tail -3; done
} > gpurun_out/r2zm_sweep.log 2>&1
cat gpurun_out/r2zm_sweep.log
