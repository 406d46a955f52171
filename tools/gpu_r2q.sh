cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
export QFLEX_KB_NOPERM=1
for v in "QFLEX_FUSED_DBG=15 QFLEX_PAIR_SA=2 QFLEX_PAIR_SC=2" "QFLEX_FUSED_DBG=15 QFLEX_PAIR_SA=4 QFLEX_PAIR_SC=2" "QFLEX_FUSED_DBG=15 QFLEX_PAIR_SA=6 QFLEX_PAIR_SC=1" "QFLEX_FUSED_DBG=15 QFLEX_PAIR_DUAL=0" "QFLEX_FUSED_DBG=15 QFLEX_PAIR_DUAL=0 QFLEX_PAIR_SA=6" "QFLEX_FUSED_DBG=13" "QFLEX_FUSED_DBG=14" "QFLEX_FUSED_DBG=11"; do
  echo "$v"; env $v python tools/kernel_bench.py 24,4,4 2>&1 | grep '"gemm"'
done > gpurun_out/r2q_loads_only.log 2>&1
for s in 22,6,6 20,8,8; do echo "dbg15 $s"; QFLEX_FUSED_DBG=15 python tools/kernel_bench.py $s 2>&1 | grep '"gemm"'; done >> gpurun_out/r2q_loads_only.log 2>&1
python - <<'PY' >> gpurun_out/r2q_loads_only.log 2>&1
import torch
x = torch.empty(1<<30, dtype=torch.float32, device='cuda')
x.fill_(1.0)
for name, fn in (("sum (read only)", lambda: x.sum()), ("fill (write only)", lambda: x.fill_(2.0)), ("copy", None)):
    if fn is None:
        y = torch.empty_like(x); fn = lambda: y.copy_(x)
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    b = 4 * (1 << 30) * (2 if name == "copy" else 1)
    print(name, "%.3f ms %.0f GB/s" % (ms, b / ms / 1e6))
PY
cat gpurun_out/r2q_loads_only.log
