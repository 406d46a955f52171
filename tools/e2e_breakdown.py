"""Where the end-to-end (host text buffers -> amplitudes) time goes for the bench workload."""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import qflex_b200
import bench
w = bench.WORKLOADS["bristlecone_70_1-32-1"]
c, o, g = bench.load_texts(w, w["cuts"][1])
lib = qflex_b200.default_library()
lib.set_verbose(1)
for i in range(3):
    t0 = time.perf_counter()
    res, st = lib.evaluate(c, o, g, memory_limit=w["memory"])
    wall = time.perf_counter() - t0
    print(json.dumps(dict(call=i, wall_ms=1e3 * wall, upload_ms=1e3 * st["seconds_upload"],
                          flatten_ms=1e3 * st["seconds_flatten"], contract_ms=1e3 * st["seconds_contract"],
                          device_ms_contract=st["device_ms_contract"])), flush=True)
