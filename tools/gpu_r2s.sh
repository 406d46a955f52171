cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
(time python3 bench.py --gpus 1 --steps 20 --warmup 5) > gpurun_out/r2s_bench_driverlike.json 2> gpurun_out/r2s_bench_driverlike.err
(time python3 bench.py --impl reference --gpus 1 --steps 4 --warmup 1) > gpurun_out/r2s_bench_reference.json 2> gpurun_out/r2s_bench_reference.err
tail -4 gpurun_out/r2s_bench_driverlike.err gpurun_out/r2s_bench_reference.err
