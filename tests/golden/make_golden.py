"""Regenerates tests/golden/amplitudes.json from the UNMODIFIED reference
(oracle/_ref/libqflex_ref.so, built by `make -C oracle` from /root/reference).

Run in the build container only (the GPU box has no /root/reference, but the
prebuilt oracle/_ref does travel there):
    OMP_NUM_THREADS=1 python tests/golden/make_golden.py
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests.support import CASES, case_texts, reference_evaluate  # noqa: E402


def main():
    # `make_golden.py name...` regenerates only those cases and keeps the rest
    only = set(sys.argv[1:])
    path = os.path.join(ROOT, "tests", "golden", "amplitudes.json")
    out = {}
    if only and os.path.exists(path):
        with open(path) as f:
            out = json.load(f)
    for name, case in CASES.items():
        if not case.get("golden", True) or (only and name not in only):
            continue
        c, o, g = case_texts(case)
        res, secs = reference_evaluate(c, o, g, case.get("initial", ""), case.get("final", ""))
        out[name] = {
            "states": [s for s, _ in res],
            "re": [float(a.real) for _, a in res],
            "im": [float(a.imag) for _, a in res],
            "reference_seconds": secs,
        }
        print(name, len(res), "amplitudes", "%.2fs" % secs, flush=True)
    with open(path, "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
