cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x -k "several_contexts or devices_flag or small_k_long or resident_program" 2>&1 | tail -5 > gpurun_out/r2m_pytest.log
( time qflex_b200/qflex_b200.x -c config/circuits/bristlecone_70_1-24-1_0.txt -o config/ordering/bristlecone_70.txt -g config/grid/bristlecone_70.txt --memory=16GB --devices=all -v 1 ) > gpurun_out/r2m_cli_alldevices.log 2>&1
( time qflex_b200/qflex_b200.x -c config/circuits/bristlecone_70_1-24-1_0.txt -o config/ordering/bristlecone_70.txt -g config/grid/bristlecone_70.txt --memory=16GB -v 1 ) > gpurun_out/r2m_cli_onedevice.log 2>&1
( time QFLEX_DEVICES=all tests/_build/refsrc/qflex_refmain.x config/circuits/bristlecone_70_1-24-1_0.txt config/ordering/bristlecone_70.txt config/grid/bristlecone_70.txt "" "" 1 16GB ) > gpurun_out/r2m_refmain_alldevices.log 2>&1
timeout 600 python bench.py --workload sycamore_53_16 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2m_bench_m16.json 2> gpurun_out/r2m_bench_m16.err
tail -5 gpurun_out/r2m_pytest.log; tail -8 gpurun_out/r2m_cli_alldevices.log; tail -6 gpurun_out/r2m_cli_onedevice.log; tail -6 gpurun_out/r2m_refmain_alldevices.log
