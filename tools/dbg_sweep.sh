# timing experiments: parts of the tensor-core kernel disabled one at a time
# (QFLEX_FUSED_DBG bits: 1 no convert, 2 no C stores, 4 no MMA, 8 no TMEM loads)
export QFLEX_KB_NOPERM=1
SHAPES=${SHAPES:-20,8,8}
for d in 0 1 2 4 8 3 6 7 10 15; do echo "dbg=$d"; QFLEX_FUSED_DBG=$d python tools/kernel_bench.py $SHAPES 2>&1 | grep gemm; done > gpurun_out/dbg_sweep.log 2>&1
