import json, os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1:
    import numpy as np
    import qflex_b200
    from tests import support
    lib = qflex_b200.default_library()
    base = dict(circuit="bristlecone_70_1-32-1_0.txt", ordering="bristlecone_70.txt", grid="bristlecone_70.txt")
    for v in ([[0],[0],[0],[0]], [[2],[9],[4],[15]], [[2],[9],[4],[0]]):
        c, o, g = support.case_texts(dict(base, cut_values=v))
        for rep in range(2):
            res, st = lib.evaluate(c, o, g, memory_limit=64 << 30, precision="auto")
            print(os.environ.get("QFLEX_TC_DEBIAS"), v, rep, repr(res[0][1]), int(st["launches_gemm_tc"]), flush=True)
else:
    for d in ("0", "1", "0"):
        subprocess.run([sys.executable, __file__, "x"], env=dict(os.environ, QFLEX_TC_DEBIAS=d))
