cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
export QFLEX_KB_NOPERM=1
for dbg in 0 32 15 47; do echo "dbg=$dbg"; QFLEX_FUSED_DBG=$dbg python tools/kernel_bench.py 24,4,4 2>&1 | grep '"gemm"'; done > gpurun_out/r2u.log 2>&1
cat gpurun_out/r2u.log
