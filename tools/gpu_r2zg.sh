cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
export QFLEX_KB_NOPERM=1
{
for gsz in 1 2 4 8 16 32; do echo "ngroup=$gsz"; QFLEX_TC_NGROUP=$gsz timeout 300 python tools/kernel_bench.py 14,14,14 15,15,15 18,12,10 20,9,8 2>&1 | grep -E '"gemm"'; done
} > gpurun_out/r2zg_sweep.log 2>&1
cat gpurun_out/r2zg_sweep.log
