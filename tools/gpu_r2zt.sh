cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 900 python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/r2zt_reference.json 2> gpurun_out/r2zt_reference.err; echo "reference arm rc=$?"; cut -c1-400 gpurun_out/r2zt_reference.json
