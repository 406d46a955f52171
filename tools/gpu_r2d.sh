cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
QFLEX_TC_DEBIAS=0 python tools/tc_bias.py > gpurun_out/r2d_tc_bias_off.log 2>&1
python tools/tc_bias.py > gpurun_out/r2d_tc_bias_on.log 2>&1
python tools/gpu_r2b.py > gpurun_out/r2d_precision.log 2>&1
python -m pytest tests -m gpu -q -x tests/test_make_ordering.py tests/test_reference_frontends.py tests/test_cli.py 2>&1 | tail -15 > gpurun_out/r2d_pytest.log
cat gpurun_out/r2d_precision.log; tail -5 gpurun_out/r2d_pytest.log
