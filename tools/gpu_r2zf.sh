cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
export QFLEX_KB_NOPERM=1
timeout 600 python -m pytest tests/test_gpu_tensorcore.py -m gpu -q -x -k "long_k_wide_n" 2>&1 | tail -3
{
for e in 4 5; do echo "mink=$e"; QFLEX_TC_PAIRLONG_MINK=$e timeout 300 python tools/kernel_bench.py 16,12,4 18,9,4 20,10,4 2>&1 | grep -E '"gemm"|"mixed"'; done
} > gpurun_out/r2zf_sweep.log 2>&1
cat gpurun_out/r2zf_sweep.log
