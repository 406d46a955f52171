cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 300 python tools/e2e_breakdown.py bristlecone_48_1-24-1 sycamore_53_8_batch64 bristlecone_70_1-24-1 > gpurun_out/r2zn_e2e_breakdown.log 2>&1; cat gpurun_out/r2zn_e2e_breakdown.log
