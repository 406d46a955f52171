export QFLEX_KB_NOPERM=1
for d in 0 1 2 4 8 3 5 6 7 9 12 15; do echo "dbg=$d"; QFLEX_FUSED_DBG=$d python tools/kernel_bench.py 20,8,8 24,4,4 12,8,16 2>&1 | grep gemm; done > gpurun_out/dbg_sweep.log 2>&1
bash tools/ncu_kernel.sh fused_20_8_8 tc_fused 20,8,8
python tools/ncu_pick.py gpurun_out/ncu_fused_20_8_8.csv > gpurun_out/ncu_fused_20_8_8_pick.txt 2>&1
