cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
export QFLEX_KB_NOPERM=1
timeout 2400 python -m pytest tests -m gpu -q -x > gpurun_out/r2ze_gpu_suite.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r2ze_gpu_suite.log
{
for c in 4 8 16; do echo "chunk=$c"; QFLEX_TC_CHUNK=$c timeout 300 python tools/kernel_bench.py 14,14,14 18,12,10 2>&1 | grep -E '"gemm"|"mixed"'; 
  QFLEX_TC_CHUNK=$c timeout 300 python -m pytest tests/test_gpu_tensorcore.py -m gpu -q -x -k "long_k_wide_n or split_k" 2>&1 | tail -2; done
} > gpurun_out/r2ze_chunk.log 2>&1
cat gpurun_out/r2ze_chunk.log
