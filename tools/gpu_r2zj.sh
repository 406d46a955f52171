cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
for w in rectangular_7x7_1-40-1 sycamore_53_14 sycamore_53_12; do
  timeout 1200 python bench.py --workload $w --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r2zj_bench_$w.json 2> gpurun_out/r2zj_bench_$w.err; echo "$w rc=$?"; cut -c1-150 gpurun_out/r2zj_bench_$w.json; tail -2 gpurun_out/r2zj_bench_$w.err
done
