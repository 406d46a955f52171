"""Where the end-to-end (host text buffers -> amplitudes) time goes: python tools/e2e_breakdown.py [workload ...]"""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import qflex_b200
import bench
names = sys.argv[1:] or ["bristlecone_70_1-32-1"]
lib = qflex_b200.default_library()
for name in names:
    w = bench.WORKLOADS[name]
    c, o, g = bench.load_texts(w, bench.workload_cuts(w, 1))
    for i in range(6):
        t0 = time.perf_counter()
        res, st = lib.evaluate(c, o, g, memory_limit=w["memory"])
        wall = time.perf_counter() - t0
        print(json.dumps(dict(workload=name, call=i, wall_ms=round(1e3 * wall, 3),
                              **{k: round(1e3 * st[k], 3) for k in st if k.startswith("seconds_")},
                              device_ms_contract=round(st["device_ms_contract"], 3),
                              launches=int(st["launches"]))), flush=True)
