cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 900 python bench.py --workload bristlecone_70_1-32-1_weak --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2g_bench_weak.json 2> gpurun_out/r2g_bench_weak.err || exit 1
bash tools/launch_list.sh > gpurun_out/r2g_launch_list.log 2>&1
mv gpurun_out/launches_summary.txt gpurun_out/r2g_launches_summary.txt
python tools/kernel_bench.py 20,8,8 24,4,4 16,8,8 12,8,16 16,12,4 20,4,4 24,3,3 > gpurun_out/r2g_kernel_bench.log 2>&1
bash tools/ncu_kernel.sh r2g_tc_pair_20_8_8 tc_pair 20,8,8
bash tools/ncu_kernel.sh r2g_tc_pair_24_4_4 tc_pair 24,4,4
python tools/ncu_pick.py gpurun_out/ncu_r2g_tc_pair_20_8_8.csv > gpurun_out/r2g_ncu_tc_pair_20_8_8.txt
python tools/ncu_pick.py gpurun_out/ncu_r2g_tc_pair_24_4_4.csv > gpurun_out/r2g_ncu_tc_pair_24_4_4.txt
rm -f gpurun_out/ncu_r2g_*_src.csv
python -m pytest tests -m gpu -q -x tests/test_gpu_e2e.py -k "several_contexts or batch_of_final" -s 2>&1 | tail -8 > gpurun_out/r2g_pytest.log
python -m pytest tests -m gpu -q -x tests/test_cli.py tests/test_make_ordering.py 2>&1 | tail -5 >> gpurun_out/r2g_pytest.log
cat gpurun_out/r2g_launches_summary.txt gpurun_out/r2g_kernel_bench.log; tail -12 gpurun_out/r2g_pytest.log
