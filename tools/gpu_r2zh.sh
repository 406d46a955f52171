cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2zh_bench_2gpu.json 2> gpurun_out/r2zh_bench_2gpu.err; echo "2gpu rc=$?"; tail -1 gpurun_out/r2zh_bench_2gpu.json | cut -c1-300
