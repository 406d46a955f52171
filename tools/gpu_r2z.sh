cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
export QFLEX_KB_NOPERM=1
timeout 900 python -m pytest tests/test_gpu_tensorcore.py tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -2
{
for d in 0 128; do echo "dbg=$d"; QFLEX_FUSED_DBG=$d timeout 120 python tools/kernel_bench.py 24,4,4 20,4,4 20,5,4 20,6,4 2>&1 | grep -E '"gemm"'; done
} > gpurun_out/r2za_sweep.log 2>&1
cat gpurun_out/r2za_sweep.log
