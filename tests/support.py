"""Shared helpers for the test-suite, bench.py's baseline legs and smoke().

Holds the list of (circuit, ordering, grid) cases, access to the compiled
reference (oracle/_ref, test infrastructure) and the CPU-emulated host library.
"""
import ctypes
import os
import re
import sys

import numpy as np

ROOT_ = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT_ not in sys.path:
    sys.path.insert(0, ROOT_)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CONFIG = os.path.join(ROOT, "config")

SIMPLE_CIRCUIT = """5
0 h 0
0 h 1
0 h 2
0 h 3
0 h 5
1 t 0
1 t 1
1 t 2
1 t 3
1 t 5
2 cz 0 1
3 cx 0 2
4 cx 1 3
5 cz 2 3
6 cz 3 5
11 cz 0 1
12 cx 0 2
13 cx 1 3
14 cz 2 3
15 cx 3 5
17 h 0
17 h 1
17 h 2
17 h 3
17 h 5"""
SIMPLE_ORDERING = """#
cut () 1 3
expand a 1
expand a 0
expand a 2
cut () 5
expand b 5
expand b 3
merge a b
"""
SIMPLE_GRID = "1 1\n1 1\n0 1\n"


from qflex_b200._host import restrict_cuts  # noqa: E402,F401  (slice subsets)


# name -> files under config/ (or inline texts), optional final state
CASES = {
    "test": dict(circuit="test.txt", ordering="test.txt", grid="test.txt"),
    "simple_1100x": dict(circuit_text=SIMPLE_CIRCUIT, ordering_text=SIMPLE_ORDERING,
                         grid_text=SIMPLE_GRID, initial="00000", final="1100x"),
    "rect2x2_0000": dict(circuit="rectangular_2x2_1-2-1_0.txt", ordering="rectangular_2x2.txt",
                         grid="rectangular_2x2.txt", final="0000"),
    "rect2x2_0110": dict(circuit="rectangular_2x2_1-2-1_0.txt", ordering="rectangular_2x2.txt",
                         grid="rectangular_2x2.txt", final="0110"),
    # no cut at all: nothing to deal over ranks (executor.cpp rank shares)
    "rect2x2_cutfree_0110": dict(
        circuit="rectangular_2x2_1-2-1_0.txt", grid="rectangular_2x2.txt", final="0110",
        ordering_text="expand A 0\nexpand A 1\nexpand B 2\nexpand B 3\nmerge A B\n"),
    "sycamore_53_4": dict(circuit="sycamore_53_4_0.txt", ordering="sycamore_53.txt",
                          grid="sycamore_53.txt"),
    "bristlecone_70_1-12-1": dict(circuit="bristlecone_70_1-12-1_0.txt",
                                  ordering="bristlecone_70.txt", grid="bristlecone_70.txt"),
    "bristlecone_70_1-16-1": dict(circuit="bristlecone_70_1-16-1_0.txt",
                                  ordering="bristlecone_70.txt", grid="bristlecone_70.txt"),
    "bristlecone_48_1-16-1": dict(circuit="bristlecone_48_1-16-1_0.txt",
                                  ordering="bristlecone_48.txt", grid="bristlecone_48.txt"),
    "bristlecone_48_1-24-1": dict(circuit="bristlecone_48_1-24-1_0.txt",
                                  ordering="bristlecone_48.txt", grid="bristlecone_48.txt"),
    # BASELINE configs[3]: 7x7 grid with the reference's ExampleOrdering schedule
    # (config/ordering/rectangular_7x7_v2.txt); 1 of the 256 slices, 4 of the 64
    # terminal-cut amplitudes (2^12x2^12x2^12 merges, 2^16x256x256 expands)
    "rect7x7_1-32-1_4amps": dict(
        circuit="rectangular_7x7_1-32-1_0.txt", ordering="rectangular_7x7_v2.txt",
        grid="rectangular_7x7.txt",
        cut_values=[[5], [9], [0, 1], [1], [0], [0, 1], [1], [0]]),
    # BASELINE configs[2] at the depth the reference's data model can hold (SURVEY
    # section 0.4: m >= 12 is not representable): Sycamore-53 m=8, the six terminal
    # cuts opened -> a batch of 64 output bitstrings per call; 1 and 2 of the 4096 slices
    "sycamore_53_8_64amps": dict(
        circuit="sycamore_53_8_0.txt", ordering="sycamore_53.txt", grid="sycamore_53.txt",
        cut_values=[[3], [7], [11]] + [[0, 1]] * 6),
    "sycamore_53_8_2slices_64amps": dict(
        circuit="sycamore_53_8_0.txt", ordering="sycamore_53.txt", grid="sycamore_53.txt",
        cut_values=[[3, 12], [7], [11]] + [[0, 1]] * 6),
    # BASELINE configs[4] stand-in: Sycamore-53 m=10, one of its 16384
    # slices (2^18 x 2^10 x 2^10 contractions; 199 s and 32.6 GB on the reference)
    "sycamore_53_10_1slice": dict(
        circuit="sycamore_53_10_0.txt", ordering="sycamore_53.txt", grid="sycamore_53.txt",
        cut_values=[[3], [7], [11]]),
    # BASELINE configs[1] itself (the bench headline): bristlecone_70 1-32-1, one and two
    # of its 65536 slices (2^20 x 256 x 256 expands, 2^28-element patches; 6.3 GB and
    # ~14-60 s per slice on the reference)
    "bristlecone_70_1-32-1_1slice": dict(
        circuit="bristlecone_70_1-32-1_0.txt", ordering="bristlecone_70.txt",
        grid="bristlecone_70.txt", cut_values=[[0], [0], [0], [0]]),
    "bristlecone_70_1-32-1_2slices": dict(
        circuit="bristlecone_70_1-32-1_0.txt", ordering="bristlecone_70.txt",
        grid="bristlecone_70.txt", cut_values=[[2], [9], [4], [0, 15]]),
    # 16 of the 4096 slices of config[1]'s little brother
    "bristlecone_70_1-24-1_16slices": dict(
        circuit="bristlecone_70_1-24-1_0.txt", ordering="bristlecone_70.txt",
        grid="bristlecone_70.txt", cut_values=[[0, 1], [0, 1], [0, 1], [0, 1]]),
}


def _read(kind, name):
    with open(os.path.join(CONFIG, kind, name)) as f:
        return f.read()


def case_texts(case):
    c = case.get("circuit_text") or _read("circuits", case["circuit"])
    o = case.get("ordering_text") or _read("ordering", case["ordering"])
    g = case.get("grid_text") or _read("grid", case["grid"])
    if case.get("cut_values"):
        o = restrict_cuts(o, case["cut_values"])
    return c, o, g


from oracle.ref_run import (REF_LIB, reference_available, reference_library,  # noqa: E402,F401
                            reference_set_threads, reference_evaluate, reference_multiply,
                            reference_reorder)


def emulated_host_library():
    """Host layer linked against oracle/device_emul.cpp (CPU, tests only)."""
    import sys
    sys.path.insert(0, ROOT)
    import __graft_entry__ as ge
    from qflex_b200._host import HostLibrary
    return HostLibrary(ge.build_emulated_host())


def rel_err(got, want):
    got, want = np.asarray(got), np.asarray(want)
    return float(np.linalg.norm(got - want) / max(np.linalg.norm(want), 1e-300))
