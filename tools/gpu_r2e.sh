cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
python tools/gpu_r2e.py > gpurun_out/r2e_determinism.log 2>&1
cat gpurun_out/r2e_determinism.log
