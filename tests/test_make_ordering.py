"""tools/make_ordering.py: orderings generated from circuit + grid alone
-- greedy bond contraction, cut search under a max-tensor constraint, fidelity widths --
must be valid qFlex orderings and reproduce the reference's amplitudes of the FULL
circuit (goldens made with the shipped orderings, whose cuts are different)."""
import importlib.util
import os
import re

import numpy as np
import pytest

from tests import support

TOL = 1e-4


def _tool():
    spec = importlib.util.spec_from_file_location(
        "make_ordering", os.path.join(support.ROOT, "tools", "make_ordering.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _terminal_cuts(ordering_text):
    """[(qubit, values)] of the shipped ordering's terminal cuts, in order: the generated
    ordering must enumerate the same output bitstrings."""
    out = []
    for line in ordering_text.splitlines():
        m = re.match(r"cut\s*\(([0-9,\s]*)\)\s+(\d+)\s*$", line.split("#")[0].strip())
        if m:
            out.append((int(m.group(2)), tuple(int(v) for v in m.group(1).split(",") if v.strip())))
    return out


def _amps(golden, name):
    g = golden[name]
    return g["states"], np.array(g["re"]) + 1j * np.array(g["im"])


# (case, max_log2 limit or None, max_cuts): full-circuit goldens only
CPU_RUNS = [("sycamore_53_4", None, 0), ("sycamore_53_4", None, 2), ("sycamore_53_4", 10, 0),
            ("sycamore_53_5", 12, 0), ("bristlecone_70_1-12-1", None, 0),
            ("bristlecone_70_1-12-1", 10, 0), ("rect2x2_0110", None, 2), ("test", None, 1),
            ("simple_1100x", None, 1)]


@pytest.mark.parametrize("name,limit,max_cuts", CPU_RUNS)
def test_generated_ordering_reproduces_the_golden_on_the_emulated_abi(emul, golden, name, limit,
                                                                      max_cuts):
    tool = _tool()
    case = support.CASES[name]
    c, o, g = support.case_texts(case)
    o2, info = tool.make_ordering(c, g, max_log2=limit, max_cuts=max_cuts,
                                  outputs=_terminal_cuts(o))
    if limit is not None:
        assert info["max_log2"] <= limit and info["cuts"]
    res, st = emul.evaluate(c, o2, g, case.get("initial", ""), case.get("final", ""),
                            memory_limit=8 << 30)
    states, want = _amps(golden, name)
    assert [s for s, _ in res] == states
    assert support.rel_err([a for _, a in res], want) < TOL
    # the tool's work model (sum over contractions of 2^(rank_a + rank_b - shared)) is
    # the planner's flop count up to the open output indices
    assert st["flops_job"] <= 1.5 * info["flops_per_slice"] * info["slices"] * 2 ** len(states)


def test_fidelity_keeps_a_prefix_of_the_cut_values():
    tool = _tool()
    frac, widths = tool.widths_for_fidelity([16, 16, 4], 0.25)
    assert 0.25 <= frac < 0.30 and all(1 <= w <= d for w, d in zip(widths, [16, 16, 4]))
    assert abs(np.prod(widths) / (16 * 16 * 4) - frac) < 1e-12
    assert tool.widths_for_fidelity([16, 4], 1.0) == (1.0, [16, 4])
    c, o, g = support.case_texts(support.CASES["sycamore_53_4"])
    text, info = tool.make_ordering(c, g, max_log2=10, fidelity=0.1)
    cut_lines = [l for l in text.splitlines() if l.startswith("cut")]
    assert len(cut_lines) == len(info["cuts"]) and info["fidelity"] >= 0.1
    assert any(re.match(r"cut \((\d+,)*\d+\) ", l) for l in cut_lines)  # restricted value lists


def test_generated_orderings_for_the_baseline_grids_plan(emul, monkeypatch):
    """rectangular_7x7 and sycamore_53 from circuit + grid alone: the planner accepts the
    ordering (every site expanded once, cuts before use, same output bitstrings) and the
    flops of the whole job are compared with the shipped ordering's.  The generator puts
    the terminal cuts on top, so its schedules count on the engine's open outputs; the
    7x7 case is only planned with projected leaves (its open-output arena is too large
    for the CPU-emulated arena of this test)."""
    tool = _tool()
    monkeypatch.setenv("QFLEX_FUSE_BITS", "0")
    for circuit, grid, shipped, limit, open_outputs in (
            ("sycamore_53_8_0.txt", "sycamore_53.txt", "sycamore_53.txt", 20, True),
            ("rectangular_7x7_1-32-1_0.txt", "rectangular_7x7.txt", "rectangular_7x7_v2.txt", 26, False)):
        if open_outputs:
            monkeypatch.delenv("QFLEX_OPEN_OUTPUTS", raising=False)
        else:
            monkeypatch.setenv("QFLEX_OPEN_OUTPUTS", "0")
        case = dict(circuit=circuit, grid=grid, ordering=shipped)
        c, o, g = support.case_texts(case)
        o2, info = tool.make_ordering(c, g, max_log2=limit, outputs=_terminal_cuts(o))
        assert info["max_log2"] <= limit
        mine = emul.prepare(c, o2, g, memory_limit=6 << 30)
        theirs = emul.prepare(c, o, g, memory_limit=6 << 30)
        assert mine.states == theirs.states
        a, b = mine.plan_stats(), theirs.plan_stats()
        if open_outputs:
            assert a["flops_job"] < b["flops_job"], (circuit, a["flops_job"], b["flops_job"])
        mine.close()
        theirs.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name,limit", [("rect7x7_16cycles_full", 16), ("rect7x7_16cycles_full", None),
                                        ("sycamore_53_5", 10), ("bristlecone_70_1-16-1", 12)])
def test_generated_ordering_reproduces_the_golden_on_the_gpu(gpu_lib, golden, name, limit):
    tool = _tool()
    case = support.CASES[name]
    c, o, g = support.case_texts(case)
    o2, info = tool.make_ordering(c, g, max_log2=limit, max_cuts=0, outputs=_terminal_cuts(o))
    res, st = gpu_lib.evaluate(c, o2, g, case.get("initial", ""), case.get("final", ""),
                               memory_limit=64 << 30)
    states, want = _amps(golden, name)
    assert [s for s, _ in res] == states
    assert support.rel_err([a for _, a in res], want) < TOL


@pytest.mark.gpu
def test_generated_and_shipped_orderings_agree_on_sycamore_m8_in_full(gpu_lib):
    """All 4096 slices x 64 bitstrings of Sycamore-53 m=8 (out of the reference's reach in
    full: ~3 h): two entirely different schedules -- the shipped ordering and a generated
    one with a 2^24 tensor limit -- must give the same 64 amplitudes."""
    tool = _tool()
    case = dict(circuit="sycamore_53_8_0.txt", ordering="sycamore_53.txt", grid="sycamore_53.txt",
                cut_values=[None, None, None] + [[0, 1]] * 6)
    c, o, g = support.case_texts(case)
    o2, info = tool.make_ordering(c, g, max_log2=24, outputs=_terminal_cuts(o))
    a, sa = gpu_lib.evaluate(c, o, g, memory_limit=64 << 30)
    b, sb = gpu_lib.evaluate(c, o2, g, memory_limit=64 << 30)
    assert [s for s, _ in a] == [s for s, _ in b] and len(a) == 64
    assert support.rel_err([v for _, v in b], [v for _, v in a]) < TOL
    assert sa["leaves"] == 4096 * 64
