cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
export QFLEX_KB_NOPERM=1
timeout 600 python -m pytest tests/test_gpu_tensorcore.py tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -2
{
for e in 1 0 1 0; do echo "pdl=$e"; QFLEX_PDL=$e timeout 120 python tools/kernel_bench.py 15,8,8 20,8,8 24,4,4 20,4,4 17,6,6 2>&1 | grep -E '"gemm"'; done
} > gpurun_out/r2zo_sweep.log 2>&1
cat gpurun_out/r2zo_sweep.log
