"""Three-way deviation (reference / FFMA / tensor-core path) on slices of bristlecone_70 1-32-1."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import qflex_b200
from tests import support
lib = qflex_b200.default_library()
base = dict(circuit="bristlecone_70_1-32-1_0.txt", ordering="bristlecone_70.txt", grid="bristlecone_70.txt")
ref = {"[[2], [9], [4], [0]]": complex(-1.487494887464011e-13, -8.805415706476805e-15),
       "[[2], [9], [4], [15]]": complex(4.652933244318584e-15, -2.2669166484289363e-14),
       "[[0], [0], [0], [0]]": complex(3.832168821637713e-14, 5.824832125965115e-14)}
for key, want in ref.items():
    v = json.loads(key)
    c, o, g = support.case_texts(dict(base, cut_values=v))
    out = {}
    for prec in ("ffma", "auto"):
        res, st = lib.evaluate(c, o, g, memory_limit=64 << 30, precision=prec)
        out[prec] = res[0][1]
    f, t = out["ffma"], out["auto"]
    print(key, "ref", want, "\n   ffma", f, "tc", t,
          "\n   |ffma-ref|/|ref| %.2e  |tc-ref|/|ref| %.2e  |tc-ffma|/|ref| %.2e" % (
              abs(f - want) / abs(want), abs(t - want) / abs(want), abs(t - f) / abs(want)), flush=True)
