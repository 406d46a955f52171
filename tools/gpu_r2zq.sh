cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
export QFLEX_KB_NOPERM=1
timeout 900 python -m pytest tests/test_gpu_tensorcore.py -m gpu -q -x 2>&1 | tail -3
{
for e in 1 0 1 0; do echo "pairlong splitk=$e"; QFLEX_TC_PAIRLONG_SPLITK=$e timeout 300 python tools/kernel_bench.py 12,8,16 12,8,14 10,10,16 2>&1 | grep -E '"gemm"'; done
} > gpurun_out/r2zq_sweep.log 2>&1
cat gpurun_out/r2zq_sweep.log
