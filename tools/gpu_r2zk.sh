cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
export QFLEX_KB_NOPERM=1
{
for e in 0 2; do echo "tail mode=$e"; QFLEX_TC_PAIRLONG_TAIL=$e timeout 300 python tools/kernel_bench.py 15,8,8 16,8,8 17,8,8 18,8,8 16,8,7 16,8,6 2>&1 | grep -E '"gemm"'; done
} > gpurun_out/r2zk_sweep.log 2>&1
cat gpurun_out/r2zk_sweep.log
