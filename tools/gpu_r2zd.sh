cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
for e in 1 0; do
  QFLEX_TC_PAIRLONG=$e timeout 900 python bench.py --workload rectangular_7x7_1-32-1 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2zd_7x7_pl$e.json 2> gpurun_out/r2zd_7x7_pl$e.err; echo "7x7 pairlong=$e rc=$?"; cut -c1-140 gpurun_out/r2zd_7x7_pl$e.json
done
for e in 1 0; do
  QFLEX_TC_PAIRLONG=$e timeout 900 python bench.py --workload sycamore_53_10 --steps 2 --warmup 2 --no-cpu-baseline > gpurun_out/r2zd_m10_pl$e.json 2> gpurun_out/r2zd_m10_pl$e.err; echo "m10 pairlong=$e rc=$?"; cut -c1-140 gpurun_out/r2zd_m10_pl$e.json
done
