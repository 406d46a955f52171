"""End-to-end parity on the B200: EvaluateCircuit through the C ABI against the
reference's golden amplitudes (tests/golden/amplitudes.json, generated from the
unmodified reference by tests/golden/make_golden.py) and, when the prebuilt
oracle/_ref travelled to the box, against the reference run live."""
import numpy as np
import pytest

from tests import support

pytestmark = pytest.mark.gpu

TOL = 1e-4  # BASELINE.json: per-run normwise relative error of the amplitude vector


def _trim_parked_buffers():
    """Arena / workspace of the previous evaluation stay parked for reuse; the large
    cases below need that memory back before asking how much is free."""
    from qflex_b200 import _device
    _device.load().qfx_trim_cache(0)


def _amps(golden, name):
    g = golden[name]
    return g["states"], np.array(g["re"]) + 1j * np.array(g["im"])


@pytest.mark.parametrize("name", list(support.CASES))
def test_amplitudes_match_reference_golden(gpu_lib, golden, name):
    case = support.CASES[name]
    c, o, g = support.case_texts(case)
    res, stats = gpu_lib.evaluate(c, o, g, case.get("initial", ""), case.get("final", ""),
                                  memory_limit=64 << 30)
    states, want = _amps(golden, name)
    assert [s for s, _ in res] == states
    err = support.rel_err([a for _, a in res], want)
    assert err < TOL, err
    assert stats["launches"] > 0


@pytest.mark.parametrize("name,n_open", [("bristlecone_48_1-24-1", 0), ("sycamore_53_8_64amps", 0),
                                         ("sycamore_53_8_2slices_64amps", 3),
                                         ("rect7x7_1-32-1_4amps", 1)])
def test_projected_and_partly_open_outputs_match_golden(gpu_lib, golden, monkeypatch, name, n_open):
    """The default keeps every terminal cut open; QFLEX_OPEN_OUTPUTS=0 is the
    reference's schedule (one projected leaf per bitstring), n > 0 a mix of both."""
    monkeypatch.setenv("QFLEX_OPEN_OUTPUTS", str(n_open))
    case = support.CASES[name]
    c, o, g = support.case_texts(case)
    res, stats = gpu_lib.evaluate(c, o, g, case.get("initial", ""), case.get("final", ""),
                                  memory_limit=64 << 30)
    states, want = _amps(golden, name)
    assert [s for s, _ in res] == states
    assert support.rel_err([a for _, a in res], want) < TOL
    monkeypatch.delenv("QFLEX_OPEN_OUTPUTS")
    res2, stats2 = gpu_lib.evaluate(c, o, g, case.get("initial", ""), case.get("final", ""),
                                    memory_limit=64 << 30)
    assert stats2["launches"] < stats["launches"] and stats2["leaves"] == stats["leaves"]
    assert support.rel_err([a for _, a in res2], [a for _, a in res]) < TOL


@pytest.mark.parametrize("precision", ["ffma", "bf16x3", "auto"])
@pytest.mark.parametrize("name", ["bristlecone_48_1-24-1", "sycamore_53_8_64amps"])
def test_precision_modes_agree(gpu_lib, golden, precision, name):
    case = support.CASES[name]
    c, o, g = support.case_texts(case)
    res, st = gpu_lib.evaluate(c, o, g, memory_limit=8 << 30, precision=precision)
    _, want = _amps(golden, name)
    assert support.rel_err([a for _, a in res], want) < TOL
    if precision == "ffma":
        assert st["launches_gemm_tc"] == 0
    if precision == "bf16x3" and name == "sycamore_53_8_64amps":
        assert st["launches_gemm_tc"] > 0


@pytest.mark.parametrize("precision", ["auto", "ffma"])
def test_headline_workload_one_slice_matches_reference(gpu_lib, golden, precision):
    """BASELINE configs[1] itself, the circuit bench.py times: bristlecone_70 1-32-1, one
    of its 65536 slices, against the unmodified reference's amplitude (71 s and 6.3 GB
    there), through the tensor-core path (2^20 x 256 x 256 on tc_pair) and through FFMA."""
    name = "bristlecone_70_1-32-1_1slice"
    c, o, g = support.case_texts(support.CASES[name])
    res, st = gpu_lib.evaluate(c, o, g, memory_limit=64 << 30, precision=precision)
    states, want = _amps(golden, name)
    assert [s for s, _ in res] == states
    assert support.rel_err([a for _, a in res], want) < TOL
    assert (st["launches_gemm_tc"] > 0) == (precision == "auto")


FUSED_GOLDENS = ["bristlecone_70_1-24-1_16slices", "bristlecone_70_1-16-1", "bristlecone_48_1-24-1",
                 "sycamore_53_4", "sycamore_53_8_64amps", "rect7x7_1-32-1_4amps"]


@pytest.mark.parametrize("name", FUSED_GOLDENS)
def test_sliced_and_fused_plans_match_golden(gpu_lib, golden, monkeypatch, name):
    """The default plan undoes bond-cut slicing as far as memory allows (2^f slices per
    pass); QFLEX_FUSE_BITS=0 executes the ordering slice by slice as the reference
    does.  Both against the reference's amplitudes."""
    case = support.CASES[name]
    texts = support.case_texts(case)
    args = (case.get("initial", ""), case.get("final", ""))
    states, want = _amps(golden, name)
    fused, st1 = gpu_lib.evaluate(*texts, *args, memory_limit=64 << 30)
    monkeypatch.setenv("QFLEX_FUSE_BITS", "0")
    sliced, st0 = gpu_lib.evaluate(*texts, *args, memory_limit=64 << 30)
    assert [s for s, _ in fused] == states == [s for s, _ in sliced]
    assert support.rel_err([a for _, a in fused], want) < TOL
    assert support.rel_err([a for _, a in sliced], want) < TOL
    assert st0["fused_bits"] == 0 and st0["leaves"] == st1["leaves"]
    if st1["fused_bits"] > 0:
        assert st1["launches"] < st0["launches"]


CUT_AWARE_GOLDENS = ["bristlecone_70_1-24-1_16slices", "sycamore_53_8_64amps",
                     "sycamore_53_10_1slice", "bristlecone_48_1-24-1", "rect7x7_1-32-1_4amps"]


@pytest.mark.parametrize("name", CUT_AWARE_GOLDENS)
def test_cut_aware_flatten_matches_reference_golden(gpu_lib, golden, monkeypatch, name):
    """The cut-aware flatten (leading bond cuts projected in the gate tensors, the sites
    next to them flattened inside the slice loop; what replaces src/read_circuit.cpp:
    548-689 for Sycamore m >= 14) forced on circuits the reference can run: the CUDA path
    against the reference's amplitudes, not against another kernel of this engine."""
    _trim_parked_buffers()
    monkeypatch.setenv("QFLEX_CUT_AWARE_FLATTEN", "1")
    case = support.CASES[name]
    res, st = gpu_lib.evaluate(*support.case_texts(case), case.get("initial", ""),
                               case.get("final", ""), memory_limit=64 << 30)
    states, want = _amps(golden, name)
    assert [s for s, _ in res] == states
    assert support.rel_err([a for _, a in res], want) < TOL


@pytest.mark.parametrize("name,max_log2", [("sycamore_53_8_64amps", 19), ("sycamore_53_10_1slice", 28),
                                           ("bristlecone_70_1-24-1_16slices", 18)])
def test_augmented_ordering_sums_to_the_golden(gpu_lib, golden, monkeypatch, name, max_log2):
    """tools/augment_ordering.py picks the bond cuts that make every intermediate fit
    2^max_log2 elements.  With its FIRST pick added on top of the shipped ordering and
    summed over ALL values of that extra cut, the amplitudes are still the golden's:
    slice by slice, with the slicing fused back, and with the cut-aware flatten (the
    configuration Sycamore m >= 14 runs in, where nothing else pins the amplitudes)."""
    import importlib.util
    import os
    spec = importlib.util.spec_from_file_location(
        "augment_ordering", os.path.join(support.ROOT, "tools", "augment_ordering.py"))
    tool = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(tool)
    case = support.CASES[name]
    c, o, g = support.case_texts(case)
    _, info = tool.augment(c, g, o, max_log2)
    assert info["added"], "the limit must force extra cuts"
    o2 = "cut () %d %d\n" % tuple(info["added"][0]) + o
    states, want = _amps(golden, name)
    base_leaves = None
    for cut_aware, fuse in (("0", "0"), ("0", "62"), ("1", "0")):
        _trim_parked_buffers()
        monkeypatch.setenv("QFLEX_CUT_AWARE_FLATTEN", cut_aware)
        monkeypatch.setenv("QFLEX_FUSE_BITS", fuse)
        res, st = gpu_lib.evaluate(c, o2, g, memory_limit=64 << 30)
        assert [s for s, _ in res] == states
        assert support.rel_err([a for _, a in res], want) < TOL, (cut_aware, fuse, info["added"])
        base_leaves = base_leaves or st["leaves"]
        assert st["leaves"] == base_leaves and st["leaves"] >= 2 * len(states)


@pytest.mark.parametrize("world", [2, 8])
def test_rank_shares_sum_to_the_full_amplitudes(gpu_lib, golden, world):
    case = support.CASES["bristlecone_70_1-16-1"]
    p = gpu_lib.prepare(*support.case_texts(case), memory_limit=8 << 30)
    full = p.contract(0, 1)
    total = sum(p.contract(r, world) for r in range(world))
    assert support.rel_err(total, full) < 1e-6
    _, want = _amps(golden, "bristlecone_70_1-16-1")
    assert support.rel_err(total, want) < TOL
    p.close()


def test_repeated_contract_is_deterministic(gpu_lib):
    case = support.CASES["bristlecone_48_1-24-1"]
    p = gpu_lib.prepare(*support.case_texts(case), memory_limit=8 << 30)
    a1 = p.contract().copy()
    p.reload()
    a2 = p.contract().copy()
    assert np.array_equal(a1, a2)
    p.close()


@pytest.mark.skipif(not support.reference_available(), reason="oracle/_ref did not travel")
def test_live_reference_run_matches(gpu_lib):
    # 4 of the 65536 slices of BASELINE configs[1] (bristlecone_70 1-32-1) would
    # take minutes on the CPU; use its 1-24-1 sibling, 2 slices
    case = dict(circuit="bristlecone_70_1-24-1_0.txt", ordering="bristlecone_70.txt",
                grid="bristlecone_70.txt", cut_values=[[3], [5], [1], [0, 7]])
    c, o, g = support.case_texts(case)
    want, _ = support.reference_evaluate(c, o, g)
    got, _ = gpu_lib.evaluate(c, o, g, memory_limit=8 << 30)
    assert [s for s, _ in got] == [s for s, _ in want]
    assert support.rel_err([a for _, a in got], [a for _, a in want]) < TOL


def test_linearity_over_cut_values_full_size(gpu_lib):
    # size-independent property at BASELINE configs[1] scale (bristlecone_70
    # 1-32-1, 2^28-element patches): the amplitude of a set of slices is the sum
    # of the amplitudes of its parts
    base = dict(circuit="bristlecone_70_1-32-1_0.txt", ordering="bristlecone_70.txt",
                grid="bristlecone_70.txt")
    def run(vals):
        c, o, g = support.case_texts(dict(base, cut_values=vals))
        return gpu_lib.evaluate(c, o, g, memory_limit=64 << 30)[0][0][1]
    both = run([[2], [9], [4], [0, 15]])
    one = run([[2], [9], [4], [0]])
    other = run([[2], [9], [4], [15]])
    assert abs(both - (one + other)) < 1e-5 * max(abs(one), abs(other))


def test_statevector_cases_on_gpu(gpu_lib):
    """The circuits the reference checks against cirq (tests/python/simulation_test.py),
    against the independent numpy state-vector simulator of tests/statevector.py;
    the reference's own bound, 1e-6 per amplitude."""
    import json
    import os
    from tests import statevector
    fix = json.load(open(os.path.join(support.ROOT, "tests", "golden", "simulation_fixtures.json")))
    psi = statevector.final_state_vector(fix["circuit_test"])
    rng = np.random.default_rng(3)
    for x in [int(v) for v in rng.integers(0, 2 ** 16, 6)]:
        final = bin(x)[2:].zfill(16)
        for o in ("ordering_test", "ordering_with_cuts_test"):
            res, _ = gpu_lib.evaluate(fix["circuit_test"], fix[o], fix["grid_test"], "", final)
            assert res[0][0] == final and abs(res[0][1] - psi[x]) < 1e-6
    psi4 = statevector.final_state_vector(fix["circuit_test_fsim"])
    for x in range(16):
        final = bin(x)[2:].zfill(4)
        res, _ = gpu_lib.evaluate(fix["circuit_test_fsim"], fix["ordering_2x2_test"],
                                  fix["grid_2x2_test"], "", final)
        assert abs(res[0][1] - psi4[x]) < 1e-6


def test_7x7_depth_40_tensor_core_and_ffma_paths_agree(gpu_lib):
    """BASELINE configs[3] as named (7x7, depth 1-40-1): 64 GB arena, 2^15 x 2^15 x 2^15
    complex merges.  The reference needs 72 GB of host memory and hours per slice, so
    there is no golden; the split-bf16 tensor-core path is checked against the
    independent FFMA kernels on one slice and two output bitstrings (same tolerance)."""
    import torch
    _trim_parked_buffers()
    if torch.cuda.mem_get_info(0)[0] < (80 << 30):
        pytest.skip("needs 80 GB of free device memory")
    texts = support.case_texts(dict(
        circuit="rectangular_7x7_1-40-1_0.txt", ordering="rectangular_7x7_v2.txt",
        grid="rectangular_7x7.txt", cut_values=[[0], [0]] + [[0]] * 5 + [[0, 1]]))
    tc, st_tc = gpu_lib.evaluate(*texts, memory_limit=70 << 30)
    ff, st_ff = gpu_lib.evaluate(*texts, memory_limit=70 << 30, precision="ffma")
    assert [s for s, _ in tc] == [s for s, _ in ff] and len(tc) == 2
    assert st_tc["launches_gemm_tc"] > 0 and st_ff["launches_gemm_tc"] == 0
    a, b = np.array([v for _, v in tc]), np.array([v for _, v in ff])
    assert np.linalg.norm(b) > 0
    assert support.rel_err(a, b) < TOL


def test_sycamore_m12_tensor_core_and_ffma_paths_agree(gpu_lib):
    """Sycamore-53 m=12, one slice (29.6 GB arena, 7.1e14 flop): beyond what the
    reference can hold on this host, so the tensor-core path is checked against the
    FFMA kernels."""
    import torch
    _trim_parked_buffers()
    if torch.cuda.mem_get_info(0)[0] < (40 << 30):
        pytest.skip("needs 40 GB of free device memory")
    texts = support.case_texts(dict(
        circuit="sycamore_53_12_0.txt", ordering="sycamore_53.txt", grid="sycamore_53.txt",
        cut_values=[[0], [0], [0]] + [[0]] * 5 + [[0, 1]]))
    tc, st_tc = gpu_lib.evaluate(*texts, memory_limit=60 << 30)
    ff, st_ff = gpu_lib.evaluate(*texts, memory_limit=60 << 30, precision="ffma")
    assert len(tc) == 2 and st_tc["launches_gemm_tc"] > 0 and st_ff["launches_gemm_tc"] == 0
    a, b = np.array([v for _, v in tc]), np.array([v for _, v in ff])
    assert np.linalg.norm(b) > 0
    assert support.rel_err(a, b) < TOL


def test_sycamore_m14_with_augmented_ordering(gpu_lib):
    """BASELINE configs[2] as named (m=14): runs with the cut-augmented ordering
    (tools/augment_ordering.py); tensor-core path against FFMA on one slice."""
    import torch
    _trim_parked_buffers()
    if torch.cuda.mem_get_info(0)[0] < (100 << 30):
        pytest.skip("needs 100 GB of free device memory")
    texts = support.case_texts(dict(
        circuit="sycamore_53_14_0.txt", ordering="sycamore_53_m14.txt", grid="sycamore_53.txt",
        cut_values=[[0]] * 9 + [[0]] * 5 + [[0, 1]]))
    tc, st_tc = gpu_lib.evaluate(*texts, memory_limit=120 << 30)
    ff, st_ff = gpu_lib.evaluate(*texts, memory_limit=120 << 30, precision="ffma")
    assert len(tc) == 2 and st_tc["launches_gemm_tc"] > 0 and st_ff["launches_gemm_tc"] == 0
    a, b = np.array([v for _, v in tc]), np.array([v for _, v in ff])
    assert np.linalg.norm(b) > 0
    assert support.rel_err(a, b) < TOL


def test_sycamore_m20_heavily_sliced_with_cut_aware_flatten(gpu_lib):
    """BASELINE configs[4] as named (m=20).  Grid tensors with four 2^10 bonds do not
    exist; the ordering carries 36 extra bond cuts (tools/augment_ordering.py) and the
    planner projects them in the gate tensors (cut-aware flatten, exact power-of-two
    range scaling).  One slice: a non-zero partial amplitude far below fp32's range
    (~1e-65), the same from the tensor-core and the FFMA paths."""
    import torch
    _trim_parked_buffers()
    if torch.cuda.mem_get_info(0)[0] < (140 << 30):
        pytest.skip("needs 140 GB of free device memory")
    texts = support.case_texts(dict(
        circuit="sycamore_53_20_0.txt", ordering="sycamore_53_m20.txt", grid="sycamore_53.txt",
        cut_values=[[0]] * 39 + [[0]] * 5 + [[0, 1]]))
    tc, st_tc = gpu_lib.evaluate(*texts, memory_limit=170 << 30)
    ff, st_ff = gpu_lib.evaluate(*texts, memory_limit=170 << 30, precision="ffma")
    a, b = np.array([v for _, v in tc]), np.array([v for _, v in ff])
    assert len(tc) == 2 and 0 < np.linalg.norm(b) < 1e-50
    assert support.rel_err(a, b) < TOL
    _trim_parked_buffers()


def test_sycamore_m16_cut_aware_flatten_paths_agree(gpu_lib, monkeypatch):
    """m=16 (22 extra cuts, cut-aware flatten chosen by the planner because the plain
    grid tensors do not exist): 4 slices x 2 bitstrings, tensor-core vs FFMA, and the
    rank shares of a 2-way split add up to the whole."""
    texts = support.case_texts(dict(
        circuit="sycamore_53_16_0.txt", ordering="sycamore_53_m16.txt", grid="sycamore_53.txt",
        cut_values=[[0, 1], [0, 1]] + [[0]] * 23 + [[0]] * 5 + [[0, 1]]))
    tc, st = gpu_lib.evaluate(*texts, memory_limit=64 << 30)
    ff, _ = gpu_lib.evaluate(*texts, memory_limit=64 << 30, precision="ffma")
    a, b = np.array([v for _, v in tc]), np.array([v for _, v in ff])
    assert st["leaves"] == 8 and 0 < np.linalg.norm(b) < 1e-20
    assert support.rel_err(a, b) < TOL
    parts = sum(np.array([v for _, v in gpu_lib.evaluate(*texts, memory_limit=64 << 30,
                                                         rank=r, world=2)[0]]) for r in range(2))
    assert support.rel_err(parts, a) < 1e-6
    monkeypatch.setenv("QFLEX_CUT_AWARE_FLATTEN", "0")
    with pytest.raises(Exception, match="2\\^32"):
        gpu_lib.evaluate(*texts, memory_limit=64 << 30)


def test_several_contexts_of_one_process_sum_to_the_golden(gpu_lib, golden):
    """The single-process multi-GPU path (QFLEX_DEVICES / --devices): on a one-GPU box the
    same device listed twice stands for two GPUs -- two arenas, two streams, two host
    threads, host-side sum."""
    import torch
    n = torch.cuda.device_count()
    devices = list(range(n)) if n > 1 else [0, 0]
    for name in ("bristlecone_70_1-24-1_16slices", "sycamore_53_8_2slices_64amps"):
        case = support.CASES[name]
        p = gpu_lib.prepare(*support.case_texts(case), memory_limit=16 << 30, devices=devices)
        amps = p.contract()
        states, want = _amps(golden, name)
        assert p.states == states and p.stats["lanes"] >= len(devices)
        assert support.rel_err(amps, want) < TOL
        p.close()


def test_batch_of_final_states_on_a_resident_circuit(gpu_lib):
    """f3: simulate() with a list of final states parses, plans and uploads once; each
    further bitstring rewrites the closing deltas and re-runs the ms-scale flatten.  Must
    equal one EvaluateCircuit per bitstring, and be much cheaper."""
    import time
    import qflex_b200
    case = dict(circuit="sycamore_53_8_0.txt", ordering="sycamore_53.txt", grid="sycamore_53.txt",
                cut_values=[[0, 1, 2, 3], [0], [0]] + [[0, 1]] * 6)
    c, o, g = support.case_texts(case)
    rng = np.random.default_rng(11)
    finals = ["".join(rng.choice(["0", "1"], 53)) for _ in range(24)]
    opts = dict(circuit=c.splitlines(), ordering=o.splitlines(), grid=g.splitlines(),
                memory_limit=8 << 30)
    qflex_b200.simulate(dict(opts, final_states=finals[:2]), lib=gpu_lib)  # warm
    t0 = time.perf_counter()
    batch = qflex_b200.simulate(dict(opts, final_states=finals), lib=gpu_lib)
    t_batch = time.perf_counter() - t0
    t0 = time.perf_counter()
    single = []
    for f in finals:
        single.extend(qflex_b200.simulate(dict(opts, final_state=f), lib=gpu_lib))
    t_single = time.perf_counter() - t0
    assert len(batch) == len(single) == 64 * len(finals)
    assert [s for s, _ in batch] == [s for s, _ in single]
    assert support.rel_err([a for _, a in batch], [a for _, a in single]) < 1e-6
    assert np.linalg.norm([a for _, a in batch]) > 0
    assert t_batch < t_single, (t_batch, t_single)
    print("batched simulate: %.1f ms per state vs %.1f ms" % (1e3 * t_batch / len(finals),
                                                              1e3 * t_single / len(finals)))
