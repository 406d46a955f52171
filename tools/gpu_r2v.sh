cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
export QFLEX_KB_NOPERM=1
for one in 0 1; do for dbg in 0 15; do echo "one_arrive=$one dbg=$dbg"; QFLEX_PAIR_ONE_ARRIVE=$one QFLEX_FUSED_DBG=$dbg timeout 120 python tools/kernel_bench.py 24,4,4 20,8,8 16,8,8 22,6,6 2>&1 | grep '"gemm"'; done; done > gpurun_out/r2v.log 2>&1
cat gpurun_out/r2v.log
timeout 600 python -m pytest tests -m gpu -q -x tests/test_gpu_tensorcore.py 2>&1 | tail -3
