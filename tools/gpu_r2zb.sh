cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
for w in sycamore_53_8_batch64 bristlecone_70_1-24-1 rectangular_7x7_1-32-1 bristlecone_48_1-24-1; do
  timeout 900 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2zb_bench_$w.json 2> gpurun_out/r2zb_bench_$w.err; echo "$w rc=$?"; cut -c1-160 gpurun_out/r2zb_bench_$w.json
done
timeout 1200 python bench.py --workload sycamore_53_10 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2zb_bench_sycamore_53_10.json 2> gpurun_out/r2zb_bench_sycamore_53_10.err; echo "m10 rc=$?"; cut -c1-160 gpurun_out/r2zb_bench_sycamore_53_10.json
