cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
SHAPES=24,4,4 bash tools/dbg_sweep.sh; mv gpurun_out/dbg_sweep.log gpurun_out/r2p_dbg_sweep_24_4_4.log
export QFLEX_KB_NOPERM=1
for v in "QFLEX_PAIR_SA=2" "QFLEX_PAIR_SA=3" "QFLEX_PAIR_SA=6 QFLEX_PAIR_SC=1" "QFLEX_PAIR_SC=2" "QFLEX_PAIR_SC=4 QFLEX_PAIR_SA=3" "QFLEX_PAIR_ACC=1" "QFLEX_PAIR_ACC=2" "QFLEX_PAIR_STAGGER=120" "QFLEX_PAIR_BLOCKED=1" "QFLEX_PAIR_DUAL=0"; do
  echo "$v"; env $v python tools/kernel_bench.py 24,4,4 2>&1 | grep gemm
done > gpurun_out/r2p_pair_tuning_24_4_4.log 2>&1
cat gpurun_out/r2p_dbg_sweep_24_4_4.log gpurun_out/r2p_pair_tuning_24_4_4.log
