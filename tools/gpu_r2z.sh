cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
export QFLEX_KB_NOPERM=1
timeout 900 python -m pytest tests/test_gpu_tensorcore.py tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -2
timeout 120 python tools/kernel_bench.py 20,8,8 16,8,8 24,4,4 2>&1 | grep -E '"gemm"' > gpurun_out/r2z9_kb.log; cat gpurun_out/r2z9_kb.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2z9_bench.json 2> gpurun_out/r2z9_bench.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/r2z9_bench.json
timeout 600 bash tools/ncu_kernel.sh r2z_pair_20_8_8 tc_pair 20,8,8
timeout 600 bash tools/ncu_kernel.sh r2z_pair_24_4_4 tc_pair 24,4,4
timeout 1500 bash tools/launch_list.sh | tail -25
