cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
export QFLEX_KB_NOPERM=1
{
for d in 0 1 2 4 8 15 79 64 65 66 68 72 14 13 11 7; do echo "dbg=$d"; QFLEX_FUSED_DBG=$d timeout 120 python tools/kernel_bench.py 24,4,4 2>&1 | grep '"gemm"'; done
for sa in 2 3 4 5 6; do for sc in 2 3 4; do echo "sa=$sa sc=$sc"; QFLEX_PAIR_SA=$sa QFLEX_PAIR_SC=$sc timeout 120 python tools/kernel_bench.py 24,4,4 2>&1 | grep '"gemm"'; done; done
echo "dual off"; QFLEX_PAIR_DUAL=0 timeout 120 python tools/kernel_bench.py 24,4,4 2>&1 | grep '"gemm"'
} > gpurun_out/r2y_sweep.log 2>&1
cat gpurun_out/r2y_sweep.log
# lanes heuristic: sycamore m=16
timeout 900 python bench.py --workload sycamore_53_16 --steps 2 --warmup 3 > gpurun_out/r2y_bench_m16.json 2> gpurun_out/r2y_bench_m16.err; echo "m16 rc=$?"; cut -c1-400 gpurun_out/r2y_bench_m16.json
