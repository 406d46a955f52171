cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
export QFLEX_KB_NOPERM=1
timeout 600 python -m pytest tests/test_gpu_tensorcore.py tests/test_gpu_kernels.py -m gpu -q -x 2>&1 | tail -2
{
for e in 1 0 1 0; do echo "tail one chunk=$e"; QFLEX_TC_PAIRLONG_TAIL_CHUNK=$e timeout 120 python tools/kernel_bench.py 16,8,8 16,8,7 2>&1 | grep -E '"gemm"'; done
} > gpurun_out/r2zu_sweep.log 2>&1
cat gpurun_out/r2zu_sweep.log
