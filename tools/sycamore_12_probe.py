"""One-off probe: EvaluateCircuit on a large Sycamore circuit, a few slices.
usage: sycamore_12_probe.py CIRCUIT ORDERING N_CUTS_TOTAL N_BINARY_CUTS [memory_gb]"""
import json
import sys
import time

sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
import qflex_b200
from tests import support

circuit, ordering, n_cuts, n_bin = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
mem = int(sys.argv[5]) if len(sys.argv) > 5 else 100
lib = qflex_b200.default_library()
lib.set_verbose(1)
cuts = [[0, 1]] * n_bin + [[0]] * (n_cuts - n_bin)
t = support.case_texts(dict(circuit=circuit, ordering=ordering, grid="sycamore_53.txt", cut_values=cuts))
for i in range(2):
    t0 = time.perf_counter()
    res, st = lib.evaluate(*t, memory_limit=mem << 30)
    print(json.dumps(dict(wall=time.perf_counter() - t0, flops=st["flops"], leaves=st["leaves"],
                          tf=st["flops"] / st["seconds_contract"] / 1e12,
                          contract_s=st["seconds_contract"], flatten_s=st["seconds_flatten"],
                          arena_gb=st["arena_bytes"] / 2 ** 30, launches=st["launches"], amp=str(res[0][1]))),
          flush=True)
