cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
export QFLEX_KB_NOPERM=1
for cap in 7 6 5 4 3; do for dbg in 15 0; do
  echo "cap=$cap dbg=$dbg"; QFLEX_MAP_RUN0_CAP=$cap QFLEX_FUSED_DBG=$dbg python tools/kernel_bench.py 24,4,4 20,8,8 2>&1 | grep '"gemm"'
done; done > gpurun_out/r2r_run0cap.log 2>&1
cat gpurun_out/r2r_run0cap.log
