cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -q -x tests/test_gpu_kernels.py tests/test_gpu_e2e.py -k "streaming or batch_of_final or cut_aware or m16 or m20 or m14 or golden" -s 2>&1 | tail -15) > gpurun_out/r2h_pytest.log 2>&1
for fold in 0 1; do
  QFLEX_FOLD_LAYOUT=$fold timeout 900 python bench.py --workload sycamore_53_16 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2h_bench_m16_fold${fold}.json 2> gpurun_out/r2h_bench_m16_fold${fold}.err
done
QFLEX_WORLDLINE_TREE=0 timeout 900 python bench.py --workload sycamore_53_16 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2h_bench_m16_chain.json 2> gpurun_out/r2h_bench_m16_chain.err
timeout 900 python bench.py --workload bristlecone_70_1-24-1 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2h_bench_b70_24.json 2> gpurun_out/r2h_bench_b70_24.err
QFLEX_STREAM_BITS=2 timeout 900 python bench.py --workload bristlecone_70_1-24-1 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2h_bench_b70_24_stream2.json 2> gpurun_out/r2h_bench_b70_24_stream2.err
timeout 1200 python bench.py --workload sycamore_53_20 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r2h_bench_m20.json 2> gpurun_out/r2h_bench_m20.err
tail -15 gpurun_out/r2h_pytest.log
