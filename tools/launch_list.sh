#!/bin/bash
# ncu launch list of the bench command (gpu__time_duration per launch, clocks untouched).
# The first bench.py run (no ncu) must exit 0 before this is called.
ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv \
    --log-file gpurun_out/launches_full.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline \
    > gpurun_out/launches_bench.log 2>&1
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/launches_full.csv")) if len(r) > 10]
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
ui = hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    name = r[ki].split("(")[0]
    v = float(r[vi].replace(",", ""))
    if r[ui] == "ns": v /= 1e3
    elif r[ui] == "ms": v *= 1e3
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
with open("gpurun_out/launches_summary.txt", "w") as f:
    f.write("# ncu --metrics gpu__time_duration.sum --clock-control none -c 20000: python bench.py --steps 1 --warmup 1 --no-cpu-baseline\n")
    f.write("# (default workload, 64 slices per pass, ~8000 launches per pass; the first 20000 launches = flatten + warm-up + timed + part of the profiled pass; per-launch times are cold-cache, serialised)\n")
    f.write("%-60s %8s %12s %7s\n" % ("kernel", "launches", "total_us", "share"))
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write("%-60s %8d %12.1f %6.1f%%\n" % (k[:60], a[0], a[1], 100 * a[1] / tot))
print(open("gpurun_out/launches_summary.txt").read())
PY
rm -f gpurun_out/launches_full.csv
