"""Extracts the circuit / grid / ordering texts embedded in the reference's Python
simulation test (tests/python/simulation_test.py:27-375, 498-548: the 16-qubit 4x4
circuit it checks against cirq, and the 4-qubit fSim circuit) into
tests/golden/simulation_fixtures.json.  Data only; run in the build container:
    python tests/golden/make_sim_fixtures.py
"""
import json
import os
import re

SRC = "/root/reference/tests/python/simulation_test.py"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "simulation_fixtures.json")


def main():
    text = open(SRC).read()
    out = {}
    for name in ("circuit_test", "grid_test", "ordering_test", "ordering_with_cuts_test",
                 "circuit_test_fsim", "grid_2x2_test", "ordering_2x2_test"):
        m = re.search(r"^%s = \"\"\"(.*?)\"\"\"" % name, text, flags=re.S | re.M)
        out[name] = m.group(1)
    with open(OUT, "w") as f:
        json.dump(out, f, indent=1)
    print({k: len(v) for k, v in out.items()})


if __name__ == "__main__":
    main()
