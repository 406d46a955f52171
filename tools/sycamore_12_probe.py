import sys, time, json
sys.path.insert(0, "/root/repo")
import qflex_b200
from tests import support
lib = qflex_b200.default_library()
lib.set_verbose(1)
t = support.case_texts(dict(circuit="sycamore_53_12_0.txt", ordering="sycamore_53.txt", grid="sycamore_53.txt",
                            cut_values=[[0], [0], [0]] + [[0]] * 6))
for i in range(2):
    t0 = time.perf_counter()
    res, st = lib.evaluate(*t, memory_limit=60 << 30)
    print(json.dumps(dict(wall=time.perf_counter() - t0, flops=st["flops"], tf=st["flops"] / st["seconds_contract"] / 1e12,
                          contract_s=st["seconds_contract"], flatten_s=st["seconds_flatten"], arena=st["arena_bytes"],
                          launches=st["launches"], amp=str(res[0]))), flush=True)
