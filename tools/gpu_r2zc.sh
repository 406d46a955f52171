cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
export QFLEX_KB_NOPERM=1
timeout 600 python -m pytest tests/test_gpu_tensorcore.py -m gpu -q -x -k "long_k_wide_n" 2>&1 | tail -15
{
for e in 1 0; do echo "pairlong=$e"; QFLEX_TC_PAIRLONG=$e timeout 300 python tools/kernel_bench.py 14,14,14 18,12,10 15,15,15 16,12,4 2>&1 | grep -E '"gemm"|"mixed"'; done
} > gpurun_out/r2zc_sweep.log 2>&1
cat gpurun_out/r2zc_sweep.log
