cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
N=$1
for wl in bristlecone_70_1-32-1 sycamore_53_10 sycamore_53_8_batch64; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 --workload $wl > gpurun_out/r2j_bench_${wl}_${N}gpu.json 2> gpurun_out/r2j_bench_${wl}_${N}gpu.err
tail -c 300 gpurun_out/r2j_bench_${wl}_${N}gpu.json
done
# single process driving all GPUs through the CLI (QFLEX_DEVICES): configs[1] sibling
/usr/bin/time -v qflex_b200/qflex_b200.x -c config/circuits/bristlecone_70_1-24-1_0.txt -o config/ordering/bristlecone_70.txt -g config/grid/bristlecone_70.txt --memory=16GB --devices=all -v 1 > gpurun_out/r2j_cli_alldevices.log 2>&1
