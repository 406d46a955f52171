cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
N=$1
for wl in bristlecone_70_1-32-1 sycamore_53_8_batch64; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 --workload $wl > gpurun_out/r2i_bench_${wl}_${N}gpu.json 2> gpurun_out/r2i_bench_${wl}_${N}gpu.err
tail -c 600 gpurun_out/r2i_bench_${wl}_${N}gpu.json
done
