cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
export QFLEX_KB_NOPERM=1
{
for p in 0 1; do echo "prec=$p (0 auto = tensor core, 1 FFMA)"; QFLEX_KB_PREC=$p timeout 120 python tools/kernel_bench.py 8,8,8 9,8,8 10,8,8 12,8,8 8,8,4 10,6,6 12,6,6 2>&1 | grep -E '"gemm"'; done
} > gpurun_out/r2zs_sweep.log 2>&1
cat gpurun_out/r2zs_sweep.log
