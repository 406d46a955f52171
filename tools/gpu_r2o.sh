cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2o_smoke.log 2>&1
(time timeout 3000 python -m pytest tests -m gpu -q -x 2>&1 | tail -8) > gpurun_out/r2o_pytest.log 2>&1
(time timeout 900 python bench.py --impl reference --steps 3 --warmup 3) > gpurun_out/r2o_bench_reference.json 2> gpurun_out/r2o_bench_reference.err
(time timeout 900 python bench.py --steps 3 --warmup 3) > gpurun_out/r2o_bench_default.json 2> gpurun_out/r2o_bench_default.err
cat gpurun_out/r2o_smoke.log; tail -8 gpurun_out/r2o_pytest.log; tail -4 gpurun_out/r2o_bench_reference.err; tail -4 gpurun_out/r2o_bench_default.err
