cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x tests/test_gpu_kernels.py 2>&1 | tail -5 > gpurun_out/r2n_pytest.log
timeout 600 python bench.py --workload sycamore_53_16 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2n_bench_m16.json 2> gpurun_out/r2n_bench_m16.err
timeout 600 python bench.py --workload rectangular_7x7_1-32-1 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2n_bench_7x7.json 2> gpurun_out/r2n_bench_7x7.err
QFLEX_GEMV_PERMUTE=0 timeout 600 python bench.py --workload rectangular_7x7_1-32-1 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2n_bench_7x7_noperm.json 2> gpurun_out/r2n_bench_7x7_noperm.err
tail -5 gpurun_out/r2n_pytest.log
python - <<'PY'
import json
for n in ('m16','7x7','7x7_noperm'):
    d=json.loads(open('gpurun_out/r2n_bench_%s.json'%n).read().strip().splitlines()[-1])
    print(n, 'value %.1f ms/step %.1f'%(d['value'], d['ms_per_step']))
    for k in d['kernel_classes'][:4]: print('   ', {a:(round(b,3) if isinstance(b,float) else b) for a,b in k.items()})
PY
