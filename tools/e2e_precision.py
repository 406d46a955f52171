"""Amplitude deviation of the tensor-core path from the FFMA path (which is
parity-tested against the reference) on slices of bristlecone_70."""
import os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import qflex_b200
from tests import support

lib = qflex_b200.default_library()
for circ, cuts in (("bristlecone_70_1-24-1_0.txt", [[0, 1], [0, 1], [0, 1], [0, 1]]),
                   ("bristlecone_70_1-32-1_0.txt", [[3], [7], [11], [2]]),
                   ("bristlecone_70_1-32-1_0.txt", [[0, 5], [1], [4], [9]])):
    c, o, g = support.case_texts(dict(circuit=circ, ordering="bristlecone_70.txt",
                                      grid="bristlecone_70.txt", cut_values=cuts))
    out = {}
    for prec in ("ffma", "auto"):
        res, st = lib.evaluate(c, o, g, memory_limit=64 << 30, precision=prec)
        out[prec] = np.array([a for _, a in res])
        out[prec + "_ms"] = st["device_ms_contract"]
    d = np.linalg.norm(out["auto"] - out["ffma"]) / np.linalg.norm(out["ffma"])
    print(json.dumps(dict(circuit=circ, cuts=cuts, rel_dev=float(d), ffma=str(out["ffma"]), tc=str(out["auto"]),
                          ffma_ms=out["ffma_ms"], tc_ms=out["auto_ms"])), flush=True)
