"""Host layer (parsers, planner, executor recursion, slice dealing) on the
CPU-emulated device ABI, against the reference's golden amplitudes."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

from tests import support

TOL = 1e-4  # BASELINE.json north_star: normwise relative error of the amplitude vector

CPU_CASES = ["test", "simple_1100x", "rect2x2_0000", "rect2x2_0110", "sycamore_53_4",
             "bristlecone_70_1-12-1", "bristlecone_48_1-16-1", "sycamore_53_8_64amps"]


def _amps(golden, name):
    g = golden[name]
    return g["states"], np.array(g["re"]) + 1j * np.array(g["im"])


@pytest.mark.parametrize("name", CPU_CASES)
def test_amplitudes_match_reference_golden(emul, golden, name):
    case = support.CASES[name]
    c, o, g = support.case_texts(case)
    res, stats = emul.evaluate(c, o, g, case.get("initial", ""), case.get("final", ""),
                               memory_limit=8 << 30)
    states, want = _amps(golden, name)
    assert [s for s, _ in res] == states
    assert support.rel_err([a for _, a in res], want) < TOL
    assert stats["n_amplitudes"] == len(states)


def test_simple_circuit_reference_unit_test_values(emul):
    # reference tests/src/evaluate_circuit_test.cpp:185-207
    res, _ = emul.evaluate(support.SIMPLE_CIRCUIT, support.SIMPLE_ORDERING, support.SIMPLE_GRID,
                           "00000", "1100x")
    assert [s for s, _ in res] == ["11000", "11001"]
    assert abs(res[0][1] - (0.10669 + 0.04419j)) < 1e-5
    assert abs(res[1][1] - (-0.01831 - 0.25758j)) < 1e-5


@pytest.mark.parametrize("world", [2, 3, 4])
def test_slices_dealt_to_ranks_sum_to_full_result(emul, golden, world):
    case = support.CASES["bristlecone_48_1-16-1"]
    c, o, g = support.case_texts(case)
    p = emul.prepare(c, o, g, memory_limit=8 << 30)
    full = p.contract(0, 1)
    leaves_full = p.stats["leaves"]
    parts, leaves = np.zeros_like(full), 0
    for r in range(world):
        parts += p.contract(r, world)
        leaves += p.stats["leaves"]
    assert leaves == leaves_full
    assert support.rel_err(parts, full) < 1e-6
    _, want = _amps(golden, "bristlecone_48_1-16-1")
    assert support.rel_err(parts, want) < TOL
    p.close()


def test_projected_leaves_match_open_outputs(emul, golden, monkeypatch):
    """QFLEX_OPEN_OUTPUTS=0 plans the reference's projected leaves (one dot per output
    bitstring and slice); the default keeps the output indices open.  Same amplitudes."""
    case = support.CASES["bristlecone_48_1-16-1"]
    texts = support.case_texts(case)
    res_open, st_open = emul.evaluate(*texts, memory_limit=8 << 30)
    monkeypatch.setenv("QFLEX_OPEN_OUTPUTS", "0")
    p = emul.prepare(*texts, memory_limit=8 << 30)
    assert [r for r in p.describe() if r["phase"] == "grid"][-1]["kind"] == "dot"
    p.close()
    res_proj, st_proj = emul.evaluate(*texts, memory_limit=8 << 30)
    assert [s for s, _ in res_open] == [s for s, _ in res_proj]
    a = np.array([v for _, v in res_open]); b = np.array([v for _, v in res_proj])
    assert support.rel_err(a, b) < 1e-6
    assert st_open["leaves"] == st_proj["leaves"] == 8
    assert st_open["launches"] < st_proj["launches"]
    # one of the two terminal cuts open, the other projected: slots still line up
    monkeypatch.setenv("QFLEX_OPEN_OUTPUTS", "1")
    res_mix, st_mix = emul.evaluate(*texts, memory_limit=8 << 30)
    assert [s for s, _ in res_mix] == [s for s, _ in res_proj]
    assert support.rel_err(np.array([v for _, v in res_mix]), b) < 1e-6
    assert st_open["launches"] < st_mix["launches"] < st_proj["launches"]
    _, want = _amps(golden, "bristlecone_48_1-16-1")
    assert support.rel_err(np.array([v for _, v in res_mix]), want) < TOL


def test_open_outputs_fall_back_when_memory_is_short(emul):
    # bristlecone_48 1-24-1: the open plan needs 41.2 MB, the projected one 39.7 MB
    texts = support.case_texts(support.CASES["bristlecone_48_1-24-1"])
    last = lambda p: [r for r in p.describe() if r["phase"] == "grid"][-1]["kind"]
    p = emul.prepare(*texts, memory_limit=64 << 20)
    assert last(p) == "amps"
    p.close()
    p = emul.prepare(*texts, memory_limit=40 * 10 ** 6)
    assert last(p) == "dot"
    p.close()
    with pytest.raises(Exception, match="exceeds memory limit"):
        emul.prepare(*texts, memory_limit=39 * 10 ** 6)


def test_restricting_cut_values_matches_reference_subset(emul, golden):
    # the mechanism bench.py uses to time a fixed subset of slices
    o = support.restrict_cuts("cut () 1 3\ncut (0,1) 5\n", [[1], None])
    assert o == "cut (1) 1 3\ncut (0,1) 5\n"


def test_plan_describe_and_flop_model(emul):
    case = support.CASES["bristlecone_48_1-16-1"]
    p = emul.prepare(*support.case_texts(case), memory_limit=8 << 30)
    rows = p.describe()
    grid = [r for r in rows if r["phase"] == "grid"]
    # the two terminal cuts stay open: the schedule ends in a 4-amplitude tensor
    assert grid[-1]["kind"] == "amps" and grid[-2]["kind"] == "contract"
    grid = grid[:-1]
    assert all(r["kind"] in ("contract", "dot", "scalar") for r in grid)
    # every grid tensor ends in one layout permute or a direct write
    assert any(r["kind"] == "permute" for r in rows if r["phase"] == "flatten")
    p.contract()
    flops = sum(8.0 * 2 ** (r["m_bits"] + r["n_bits"] + r["k_bits"]) for r in grid)
    assert abs(flops - p.stats["flops_per_pass"]) < 1e-6 * flops
    p.close()


@pytest.mark.parametrize("bad,needle", [
    (dict(ordering="expand a 1\nexpand a 1\n"), "is contracted multiple times"),
    # the scenarios of the reference's ValidateOrdering tests
    # (tests/src/contraction_utils_test.cpp:104-224), through the text front-end
    (dict(ordering="cut () 1 3\ncut (2,3) 1 3\n"), "Index (0,1),(1,1) is cut multiple times."),
    (dict(ordering="merge a b\nmerge a c\n"), "Patch a is merged multiple times."),
    (dict(ordering="merge a b\nexpand a 1\n"),
     "Tensor at (0,1) is added to non-empty patch a after a cut."),
    (dict(ordering="expand a 1\ncut () 1 3\nexpand a 2\n"),
     "Tensor at (1,0) is added to non-empty patch a after a cut."),
    (dict(ordering="expand b 1\ncut () 1 3\nmerge a b\n"),
     "Patch a is merged into non-empty patch b after a cut."),
    (dict(ordering="merge a b\ncut () 1 3\nexpand b 1\n"),
     "Tensor at (0,1) is added to non-empty patch b after a cut."),
    (dict(ordering="merge a c\ncut () 1 3\nmerge b c\n"),
     "Patch b is merged into non-empty patch c after a cut."),
    (dict(ordering="frobnicate a 1\n"), "Received an invalid operation in config."),
    (dict(ordering="expand a 9\n"), "Index 9 must be within grid boundaries."),
    (dict(ordering="expand a 4\n"), "Index must specify an active qubit."),
    (dict(ordering="cut (a) 1 3\n"), "Cut values must be comma-separated ints"),
    (dict(circuit="5\n0 h 0\n1 cz 0 3\n"), "are not nearest neighbors"),
    (dict(circuit="4\n0 h 0\n"), "does not match the number of active qubits"),
    (dict(circuit="5\n1 h 0\n0 h 1\n"), "Cycle number can only increase."),
    (dict(circuit="5\n0 h 0\n0 t 0\n"), "Qubits can only used once per cycle."),
    (dict(circuit="5\n0 foo 0\n"), "Invalid gate name provided: foo"),
    (dict(grid="1 1\n1\n"), "is inconsistent with a width"),
    (dict(final="11"), "must be equal to size of initial_conf"),
    (dict(ordering="expand a 1\nexpand a 0\n", final="11000"), "hand side of pair not found"),
    (dict(memory=1024), "exceeds memory limit"),
])
def test_error_messages_follow_reference(emul, bad, needle):
    from qflex_b200 import QflexError
    with pytest.raises(QflexError) as e:
        emul.evaluate(bad.get("circuit", support.SIMPLE_CIRCUIT),
                      bad.get("ordering", support.SIMPLE_ORDERING),
                      bad.get("grid", support.SIMPLE_GRID), "00000",
                      bad.get("final", "1100x"), memory_limit=bad.get("memory"))
    assert needle in str(e.value)


def test_simulate_dict_surface_prints_and_returns_empty_on_error(emul, capsys):
    import qflex_b200
    out = qflex_b200.simulate({"circuit": ["5", "0 foo 0"], "ordering": ["expand a 0"],
                               "grid": ["1 1", "1 1", "0 1"]}, lib=emul)
    assert out == []
    assert "Invalid gate name provided" in capsys.readouterr().err
    ok = qflex_b200.simulate({
        "circuit_filename": os.path.join(support.CONFIG, "circuits", "test.txt"),
        "ordering_filename": os.path.join(support.CONFIG, "ordering", "test.txt"),
        "grid_filename": os.path.join(support.CONFIG, "grid", "test.txt")}, lib=emul)
    assert [s for s, _ in ok] == ["00000", "00001"]


def test_product_libraries_export_the_declared_abi():
    """Every function declared in include/*.h is exported by the in-tree
    product libraries (no compute call: there is no GPU here)."""
    root = support.ROOT
    for header, lib in (("qflex_b200.h", "libqflex_b200.so"),
                        ("qflex_b200_host.h", "libqflex_host.so")):
        text = open(os.path.join(root, "include", header)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names = set(re.findall(r"\b(qfx_\w+|qflex_b200_\w+)\s*\(", text))
        assert names, header
        path = os.path.join(root, "qflex_b200", lib)
        assert os.path.exists(path), "run __graft_entry__.build() first"
        handle = ctypes.CDLL(path)
        for n in sorted(names):
            assert hasattr(handle, n), "%s does not export %s" % (lib, n)


def test_product_path_fails_loudly_without_gpu():
    """No CPU fallback: without a CUDA device the product library must raise."""
    import qflex_b200
    from qflex_b200 import _device
    c = _device.load()
    n = ctypes.c_int(0)
    rc = c.qfx_device_count(ctypes.byref(n))
    if rc == 0 and n.value > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(qflex_b200.QflexError):
        qflex_b200.default_library().evaluate(*support.case_texts(support.CASES["test"]))


def test_kernels_are_sm100a_sass():
    out = subprocess.run(["cuobjdump", "-lelf",
                          os.path.join(support.ROOT, "qflex_b200", "libqflex_b200.so")],
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    assert "sm_100a" in out.stdout


def test_augmented_ordering_keeps_the_amplitudes(emul, golden, monkeypatch):
    """tools/augment_ordering.py: bond cuts added until every intermediate fits; the
    planner's sizes agree with the tool's graph replay and the amplitudes (sums over
    the extra slices) do not move."""
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "augment_ordering", os.path.join(support.ROOT, "tools", "augment_ordering.py"))
    tool = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(tool)
    case = support.CASES["bristlecone_48_1-16-1"]
    c, o, g = support.case_texts(case)
    same, info0 = tool.augment(c, g, o, 60)
    assert info0["added"] == [] and info0["max_log2"] == 14
    assert [l for l in same.splitlines() if l.strip()] == [l for l in o.splitlines() if l.strip()]
    o2, info = tool.augment(c, g, o, 12)
    assert info["added"] and info["max_log2"] <= 12
    res, st = emul.evaluate(c, o2, g, memory_limit=8 << 30)
    states, want = _amps(golden, "bristlecone_48_1-16-1")
    assert [s for s, _ in res] == states
    assert support.rel_err(np.array([v for _, v in res]), want) < TOL
    res0, st0 = emul.evaluate(c, o, g, memory_limit=8 << 30)
    assert st["leaves"] == 4 * st0["leaves"]  # the cut bond has dimension 4
    # the tool's flop count per slice is the planner's (ordering as written)
    monkeypatch.setenv("QFLEX_FUSE_BITS", "0")
    p = emul.prepare(c, o, g, memory_limit=8 << 30)
    p.contract()
    assert abs(p.stats["flops_per_pass"] - info0["flops_per_slice"]) < 0.05 * info0["flops_per_slice"]  # the tool ignores output indices
    p.close()


@pytest.mark.parametrize("name", ["test", "bristlecone_48_1-16-1", "sycamore_53_4",
                                  "bristlecone_70_1-12-1", "rect2x2_0000"])
def test_cut_aware_flatten_is_bit_identical(emul, monkeypatch, name):
    """QFLEX_CUT_AWARE_FLATTEN=1: the leading bond cuts are applied to the gate tensors
    and the sites next to them are flattened inside the slice loop.  Projecting a leg
    before or after the world-line products selects the same numbers, and the range
    scaling is by powers of two, so the amplitudes are the same bits."""
    case = support.CASES[name]
    texts = support.case_texts(case)
    args = (case.get("initial", ""), case.get("final", ""))
    has_bond_cut = any(l.split("#")[0].strip().startswith("cut") and len(l.split(")")[1].split()) == 2
                       for l in texts[1].splitlines() if l.split("#")[0].strip())
    monkeypatch.setenv("QFLEX_FUSE_BITS", "0")  # slice by slice, as the ordering is written
    r0, s0 = emul.evaluate(*texts, *args, memory_limit=8 << 30)
    monkeypatch.setenv("QFLEX_CUT_AWARE_FLATTEN", "1")
    monkeypatch.setenv("QFLEX_WORLDLINE_TREE", "0")  # world lines left to right, as the reference
    r1, s1 = emul.evaluate(*texts, *args, memory_limit=8 << 30)
    assert r0 == r1
    # default: the in-loop world lines are contracted as balanced trees (same factors,
    # another association order: equal to rounding), optionally writing the grid tensor
    # in its final layout from the root
    seen = {}
    for fold, programs in (("0", "0"), ("1", "0"), ("0", "1"), ("1", "1")):
        monkeypatch.setenv("QFLEX_WORLDLINE_TREE", "1")
        monkeypatch.setenv("QFLEX_FOLD_LAYOUT", fold)
        # the small tree steps of all in-loop sites as ONE resident program per slice
        # (qfx_program_run): the same arithmetic in the same order, fewer launches
        monkeypatch.setenv("QFLEX_LINE_PROGRAMS", programs)
        r2, s2 = emul.evaluate(*texts, *args, memory_limit=8 << 30)
        assert [s for s, _ in r2] == [s for s, _ in r0]
        assert support.rel_err([a for _, a in r2], [a for _, a in r0]) < 1e-5
        assert s2["bytes"] <= s1["bytes"]
        seen[(fold, programs)] = (r2, s2)
    for fold in ("0", "1"):
        assert seen[(fold, "1")][0] == seen[(fold, "0")][0]
        if has_bond_cut:
            assert seen[(fold, "1")][1]["launches"] < seen[(fold, "0")][1]["launches"]
    has_bond_cut = any(l.split("#")[0].strip().startswith("cut") and len(l.split(")")[1].split()) == 2
                       for l in texts[1].splitlines() if l.split("#")[0].strip())
    if has_bond_cut:
        assert s1["launches"] > s0["launches"] and s1["arena_bytes"] < s0["arena_bytes"]


FUSE_CASES = ["test", "simple_1100x", "sycamore_53_4", "bristlecone_70_1-12-1",
              "bristlecone_48_1-16-1", "rect2x2_0000"]


@pytest.mark.parametrize("name", FUSE_CASES)
def test_fused_cut_bits_keep_the_amplitudes(emul, golden, monkeypatch, name):
    """The planner undoes bond-cut slicing as far as the memory limit allows (low bits
    of a cut bond stay an ordinary bond): fewer, larger launches, the same amplitudes
    and the same leaf count as the ordering executed slice by slice (QFLEX_FUSE_BITS=0)."""
    case = support.CASES[name]
    texts = support.case_texts(case)
    args = (case.get("initial", ""), case.get("final", ""))
    states, want = _amps(golden, name)
    monkeypatch.setenv("QFLEX_FUSE_BITS", "0")
    r0, s0 = emul.evaluate(*texts, *args, memory_limit=8 << 30)
    assert s0["fused_bits"] == 0
    monkeypatch.delenv("QFLEX_FUSE_BITS")
    r1, s1 = emul.evaluate(*texts, *args, memory_limit=8 << 30)
    assert [s for s, _ in r1] == states == [s for s, _ in r0]
    assert support.rel_err([a for _, a in r1], want) < TOL
    assert support.rel_err([a for _, a in r1], [a for _, a in r0]) < 1e-5
    assert s1["leaves"] == s0["leaves"]
    has_bond_cut = any(l.split("#")[0].strip().startswith("cut") and len(l.split(")")[1].split()) == 2
                       for l in texts[1].splitlines() if l.split("#")[0].strip())
    if has_bond_cut:
        assert s1["fused_bits"] > 0 and s1["launches"] < s0["launches"]
        assert s1["flops"] <= s0["flops"]
    # one bit at a time: every intermediate setting is the same sum
    monkeypatch.setenv("QFLEX_FUSE_BITS", "1")
    r2, s2 = emul.evaluate(*texts, *args, memory_limit=8 << 30)
    assert s2["fused_bits"] <= 1 and support.rel_err([a for _, a in r2], want) < TOL


def test_fusion_respects_the_memory_limit_and_the_rank_count(emul, golden):
    case = support.CASES["bristlecone_70_1-12-1"]
    c, o, g = support.case_texts(case)
    full = emul.prepare(c, o, g, memory_limit=8 << 30)
    full.contract()
    small = emul.prepare(c, o, g, memory_limit=1500000)
    small.contract()
    assert small.stats["arena_bytes"] <= 1500000
    assert small.stats["fused_bits"] < full.stats["fused_bits"]
    _, want = _amps(golden, "bristlecone_70_1-12-1")
    # planned for 8 ranks: at least 8 slices stay apart, every rank gets a share
    p8 = emul.prepare(c, o, g, memory_limit=8 << 30, world=8)
    parts, shares = np.zeros(len(want), dtype=complex), []
    for r in range(8):
        parts += p8.contract(r, 8)
        shares.append(p8.stats["leaves"])
    assert min(shares) > 0 and sum(shares) == full.stats["leaves"]
    assert p8.stats["fused_bits"] == full.stats["fused_bits"] - 3
    assert support.rel_err(parts, want) < TOL
    for p in (full, small, p8):
        p.close()


def test_unaligned_cut_values_are_not_fused(emul, golden):
    name = "sycamore_53_8_2slices_64amps"   # first cut: values (3, 12)
    case = support.CASES[name]
    res, st = emul.evaluate(*support.case_texts(case), memory_limit=8 << 30)
    assert st["fused_bits"] == 0
    _, want = _amps(golden, name)
    assert support.rel_err([a for _, a in res], want) < TOL


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("ordering", ["cut_free", "terminal_only"])
def test_rank_shares_without_a_dealt_cut(emul, world, ordering):
    """with no projected cut the accumulate steps ran on every rank and the
    all-reduce returned world x the amplitudes."""
    import json
    fix = json.load(open(os.path.join(support.ROOT, "tests", "golden", "simulation_fixtures.json")))
    c, g = fix["circuit_test_fsim"], fix["grid_2x2_test"]
    o = "\n".join(l for l in fix["ordering_2x2_test"].splitlines() if not l.startswith("cut"))
    final = "0110"
    if ordering == "terminal_only":
        o = "cut () 3\ncut () 2\n" + o
        final = "01xx"
    p = emul.prepare(c, o, g, "", final)
    full = p.contract(0, 1).copy()
    assert np.linalg.norm(full) > 0
    total = sum(p.contract(r, world) for r in range(world))
    assert support.rel_err(total, full) < 1e-12
    p.close()


@pytest.mark.parametrize("name", ["bristlecone_70_1-12-1", "bristlecone_48_1-16-1",
                                  "rect2x2_cutfree_0110", "test"])
@pytest.mark.parametrize("devices", [[0, 1], [0, 1, 2, 3, 4, 5, 6, 7]])
def test_one_process_driving_several_devices(emul, golden, name, devices):
    """Multi-GPU inside the drop-in entry points (what QFLEX_DEVICES / --devices select):
    every device holds the network, the slices are dealt over (lane, device), partial
    amplitudes are summed on the host.  On the emulated ABI a device is just a context."""
    case = support.CASES[name]
    c, o, g = support.case_texts(case)
    p = emul.prepare(c, o, g, case.get("initial", ""), case.get("final", ""),
                     memory_limit=8 << 30, devices=devices)
    amps = p.contract()
    states, want = _amps(golden, name)
    assert p.states == states
    assert support.rel_err(amps, want) < TOL
    assert p.stats["lanes"] >= len(devices)
    one = emul.prepare(c, o, g, case.get("initial", ""), case.get("final", ""), memory_limit=8 << 30)
    one.contract()
    assert p.stats["leaves"] == one.stats["leaves"]  # every leaf exactly once
    # ... and on top of a rank split: 2 ranks x these devices
    parts = sum(p.contract(r, 2) for r in range(2))
    assert support.rel_err(parts, want) < TOL
    p.close()
    one.close()


def test_device_list_from_the_environment(emul, golden, monkeypatch):
    """The reference-signature entry point EvaluateCircuit(QflexInput*) (and with it the
    unmodified reference front-ends) picks its GPUs from QFLEX_DEVICES."""
    import subprocess
    import sys
    cli = os.path.join(support.ROOT, "tests", "_build", "refsrc_emul", "qflex_refmain.x")
    if not os.path.exists(cli):
        pytest.skip("reference front-end not built (needs /root/reference)")
    cfg = support.CONFIG
    args = [cli, cfg + "/circuits/test.txt", cfg + "/ordering/test.txt", cfg + "/grid/test.txt"]
    base = subprocess.run(args, stdout=subprocess.PIPE, text=True, env=dict(os.environ)).stdout
    multi = subprocess.run(args, stdout=subprocess.PIPE, text=True,
                           env=dict(os.environ, QFLEX_DEVICES="0,1,2")).stdout
    assert base.strip() and [l.split(":")[0] for l in base.splitlines()] == \
        [l.split(":")[0] for l in multi.splitlines()]
    bad = subprocess.run(args, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                         env=dict(os.environ, QFLEX_DEVICES="0;1"))
    assert bad.returncode == 2 and "Device list" in bad.stderr
