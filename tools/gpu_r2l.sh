cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
for L in 2 4 8; do
QFLEX_LANES=$L timeout 900 python bench.py --workload sycamore_53_16 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2l_bench_m16_lanes$L.json 2> gpurun_out/r2l_bench_m16_lanes$L.err
python -c "
import json; d=json.loads(open('gpurun_out/r2l_bench_m16_lanes$L.json').read().strip().splitlines()[-1]); print('lanes $L', d['value'], d['ms_per_step'], d['e2e']['value'])"
done
python -m pytest tests -m gpu -q -x tests/test_gpu_kernels.py -k "resident_program" 2>&1 | tail -3
