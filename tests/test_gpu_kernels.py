"""Parity of the CUDA kernels (through the C ABI) with the oracle."""
import numpy as np
import pytest

from oracle import kernel_oracle as ko
from tests import support

pytestmark = pytest.mark.gpu

TOL = 1e-5  # fp32 accumulation vs the oracle evaluated in complex128, normwise


def _rand(n, rng):
    return (rng.uniform(-1, 1, n) + 1j * rng.uniform(-1, 1, n)).astype(np.complex64)


@pytest.fixture(scope="module")
def ctx():
    from qflex_b200 import _device
    c = _device.Context(0)
    yield c
    c.close()


def _random_tables(rng, m, n, k):
    """Random placement of the m/k bits inside A, k/n inside B, m/n inside C."""
    pa = rng.permutation(m + k)
    pb = rng.permutation(n + k)
    pc = rng.permutation(m + n)
    return dict(a_m=[int(x) for x in pa[:m]], a_k=[int(x) for x in pa[m:]],
                b_k=[int(x) for x in pb[:k]], b_n=[int(x) for x in pb[k:]],
                c_m=[int(x) for x in pc[:m]], c_n=[int(x) for x in pc[m:]])


def _run_contract(ctx, t, rng, off_jitter=0):
    from qflex_b200 import _device
    m, n, k = len(t["a_m"]), len(t["b_n"]), len(t["a_k"])
    sa, sb, sc = 1 << (m + k), 1 << (n + k), 1 << (m + n)
    a, b = _rand(sa, rng), _rand(sb, rng)
    oa = off_jitter
    ob = oa + sa + 32 + off_jitter
    oc = ob + sb + 32 + off_jitter
    ctx.reserve(oc + sc + 64)
    ctx.upload(oa, a)
    ctx.upload(ob, b)
    ctx.upload(oc, np.zeros(sc, dtype=np.complex64))
    ctx.contract(_device.make_desc(oa, ob, oc, t["a_m"], t["c_m"], t["b_n"], t["c_n"],
                                   t["a_k"], t["b_k"]))
    got = ctx.download(oc, sc)
    want = ko.contract(a, b, acc_dtype=np.complex128, **t)
    return support.rel_err(got, want)


# (m_bits, n_bits, k_bits): every tile configuration, ragged tiles (m < tile),
# split-K (small output, long K), tiny outputs (dot kernel), K = 1 and K = 2.
SHAPES = [(0, 0, 0), (0, 0, 5), (0, 0, 14), (3, 0, 0), (1, 1, 1), (2, 3, 9), (6, 0, 12),
          (4, 1, 1), (10, 1, 1), (11, 2, 1), (12, 3, 3), (13, 4, 4), (9, 5, 5), (12, 5, 3),
          (8, 6, 6), (7, 6, 10), (6, 6, 8), (10, 7, 6), (8, 8, 8), (7, 7, 0), (9, 7, 12),
          (3, 10, 4), (0, 9, 7), (14, 0, 2), (5, 4, 13), (15, 1, 0), (9, 9, 3)]


@pytest.mark.parametrize("m,n,k", SHAPES)
def test_contract_random_bit_placement(ctx, m, n, k):
    rng = np.random.default_rng(1000 * m + 100 * n + k)
    t = _random_tables(rng, m, n, k)
    assert _run_contract(ctx, t, rng) < TOL


@pytest.mark.parametrize("outs,k,seed", [(1, 10, 0), (1, 17, 1), (2, 12, 2), (3, 15, 3), (5, 14, 4),
                                         (6, 16, 5), (6, 10, 6), (4, 20, 7), (5, 13, 8)])
@pytest.mark.parametrize("side", ["a", "b"])
def test_contract_few_outputs_long_k_gemv(ctx, outs, k, seed, side):
    """The tail of an open-output schedule: 2..64 outputs all on one operand, long K,
    output bits anywhere inside the big operand (gemv_rows reads it once, linearly)."""
    rng = np.random.default_rng(4000 + seed)
    t = _random_tables(rng, outs, 0, k) if side == "a" else _random_tables(rng, 0, outs, k)
    assert _run_contract(ctx, t, rng) < TOL
    # output bits lowest / highest in the big operand's address
    big, free = ("a_m", "a_k") if side == "a" else ("b_n", "b_k")
    for lowest in (True, False):
        t2 = dict(t)
        t2[big] = list(range(outs)) if lowest else list(range(k, k + outs))
        t2[free] = list(range(outs, outs + k)) if lowest else list(range(k))
        assert _run_contract(ctx, t2, rng) < TOL


@pytest.mark.parametrize("m,n,k", [(16, 1, 1), (18, 2, 1), (17, 0, 2), (20, 2, 2), (16, 0, 0),
                                   (19, 1, 2), (21, 1, 0), (1, 17, 1), (2, 18, 2), (16, 3, 3),
                                   (18, 3, 2), (17, 2, 3), (20, 3, 3), (16, 0, 3), (16, 3, 0),
                                   (3, 19, 3)])
def test_contract_long_m_tiny_n_k_streaming(ctx, m, n, k):
    """World-line steps on large grid tensors (cut-aware flatten): N, K <= 4, long M."""
    rng = np.random.default_rng(77 * m + 10 * n + k)
    assert _run_contract(ctx, _random_tables(rng, m, n, k), rng) < TOL
    # the layout a flatten step produces: A = [m_hi | k | m_lo], C = [m_hi, m_lo | n]
    lo = min(m, n) if False else 5
    big, small = (m, n) if m >= n else (n, m)
    a_free = list(range(lo)) + list(range(lo + k, big + k))
    a_k = list(range(lo, lo + k))
    c_big = list(range(small, small + big))
    t = dict(a_m=a_free, a_k=a_k, b_k=list(range(small, small + k)), b_n=list(range(small)),
             c_m=c_big, c_n=list(range(small)))
    if m < n:  # same contraction with the roles of the operands exchanged
        t = dict(a_m=t["b_n"], a_k=t["b_k"], b_k=t["a_k"], b_n=t["a_m"], c_m=t["c_n"], c_n=t["c_m"])
    assert _run_contract(ctx, t, rng) < TOL
    assert _run_contract(ctx, t, rng, off_jitter=1) < TOL


@pytest.mark.parametrize("m,n,k", [(8, 6, 6), (12, 3, 3), (9, 7, 5), (0, 0, 10)])
def test_contract_odd_base_offsets(ctx, m, n, k):
    # cut projections shift operand bases; odd bases disable 16-byte accesses
    rng = np.random.default_rng(7 * m + n + k)
    t = _random_tables(rng, m, n, k)
    assert _run_contract(ctx, t, rng, off_jitter=1) < TOL


def test_contract_reference_multiply_golden(ctx):
    # reference tests/src/tensor_test.cpp:496-530: {20, 14, 100, 78}
    from qflex_b200 import _device
    a = np.arange(8).astype(np.complex64)
    b = np.arange(8, 0, -1).astype(np.complex64)
    t, idx, dims = ko.bits_from_named(["a", "b", "c"], [2, 2, 2], ["b", "c", "d"], [2, 2, 2])
    ctx.reserve(256)
    ctx.upload(0, a)
    ctx.upload(32, b)
    ctx.contract(_device.make_desc(0, 32, 64, t["a_m"], t["c_m"], t["b_n"], t["c_n"],
                                   t["a_k"], t["b_k"]))
    np.testing.assert_allclose(ctx.download(64, 4).real, [20, 14, 100, 78])
    assert idx == ["a", "d"] and dims == [2, 2]


def test_contract_named_layout_as_planner_builds_it(ctx):
    # the dominant shape family of the Bristlecone schedules: patch [.. k .. k ..]
    # times grid tensor [k, k, r, r], d = 8 (bristlecone_48 1-24-1: 2^15 x 64 x 64)
    rng = np.random.default_rng(5)
    a_idx, a_dims = ["p", "x", "q", "y", "r", "s", "t"], [8] * 7
    b_idx, b_dims = ["y", "x", "u", "v"], [8] * 4
    t, _, _ = ko.bits_from_named(a_idx, a_dims, b_idx, b_dims)
    assert (len(t["a_m"]), len(t["b_n"]), len(t["a_k"])) == (15, 6, 6)
    assert _run_contract(ctx, t, rng) < TOL


def test_contract_rejects_overlap_and_bad_tables(ctx):
    from qflex_b200 import _device
    ctx.reserve(1024)
    d = _device.make_desc(0, 64, 32, [1, 2, 3, 4, 5], [0, 1, 2, 3, 4], [], [], [0], [0])
    with pytest.raises(_device.DeviceError, match="overlaps"):
        ctx.contract(d)
    d = _device.make_desc(0, 64, 128, [1, 1], [0, 1], [], [], [0], [0])
    with pytest.raises(_device.DeviceError, match="distinct"):
        ctx.contract(d)
    d = _device.make_desc(1 << 40, 64, 128, [1, 2, 3, 4, 5], [0, 1, 2, 3, 4], [], [], [0], [0])
    with pytest.raises(_device.DeviceError, match="outside the arena"):
        ctx.contract(d)


@pytest.mark.parametrize("k", [0, 3, 8, 13, 20])
def test_dot_accumulate_complex128(ctx, k):
    from qflex_b200 import _device
    rng = np.random.default_rng(k)
    n = 1 << k
    a, b = _rand(n, rng), _rand(n, rng)
    pa, pb = [int(x) for x in rng.permutation(k)], [int(x) for x in rng.permutation(k)]
    ctx.reserve(2 * n + 64)
    ctx.upload(0, a)
    ctx.upload(n + 32, b)
    ctx.amps_reserve(3)
    d = _device.make_desc(0, n + 32, 0, [], [], [], [], pa, pb)
    ctx.dot_accumulate(d, 1)
    ctx.dot_accumulate(d, 1)  # accumulates
    amps = ctx.amps(3)
    want = ko.dot(a, b, pa, pb)
    assert amps[0] == 0 and amps[2] == 0
    assert abs(amps[1] - 2 * want) <= 1e-6 * max(1.0, np.sqrt(n))


@pytest.mark.parametrize("nbits", [0, 1, 6, 10, 14])
def test_tensor_accumulate_open_outputs(ctx, nbits):
    """qfx_tensor_accumulate: 2^k amplitudes of a final tensor added (complex128) into
    mixed-radix slots -- bit-exact: every slot receives (double) of one fp32 value."""
    from qflex_b200 import _device
    rng = np.random.default_rng(nbits)
    n = 1 << nbits
    t = _rand(n, rng)
    ctx.reserve(n + 64)
    ctx.upload(7, t)
    # slot strides: a bit permutation spread over a table 3x as large (gaps = slots of
    # projected terminal cuts), starting at slot 2
    perm = [int(x) for x in rng.permutation(nbits)]
    strides = [3 << p for p in perm]
    total = 2 + 3 * n + 1
    ctx.amps_reserve(total)
    ctx.tensor_accumulate(7, strides, 2)
    ctx.tensor_accumulate(7, strides, 2)  # accumulates
    got = ctx.amps(total)
    want = np.zeros(total, dtype=np.complex128)
    for i in range(n):
        slot = 2 + sum(strides[b] for b in range(nbits) if (i >> b) & 1)
        want[slot] += 2 * np.complex128(t[i])
    assert np.array_equal(got, want)
    if nbits >= 2:
        with pytest.raises(_device.DeviceError, match="overlap"):
            ctx.tensor_accumulate(7, [1, 2, 3] + [8 << b for b in range(nbits - 3)][:nbits - 3]
                                  if nbits >= 3 else [1, 1], 0)
    with pytest.raises(_device.DeviceError, match="outside the amplitude table"):
        ctx.tensor_accumulate(7, strides, total)


@pytest.mark.parametrize("nbits,seed", [(0, 0), (1, 0), (3, 1), (5, 2), (9, 3), (10, 4), (12, 5),
                                        (12, 6), (16, 7), (20, 8), (22, 9)])
def test_permute_bit_exact(ctx, nbits, seed):
    rng = np.random.default_rng(seed)
    n = 1 << nbits
    src = _rand(n, rng)
    dst_pos = [int(x) for x in rng.permutation(nbits)]
    ctx.reserve(2 * n + 64)
    ctx.upload(0, src)
    ctx.permute(0, n + 32, dst_pos)
    got = ctx.download(n + 32, n)
    assert np.array_equal(got, ko.permute(src, dst_pos))


def test_permute_reference_patterns(ctx):
    # right-only / left-only / worst-case L-R-L patterns of
    # tests/src/tensor_test.cpp:271-469, as bit permutations of a 2^16 tensor
    rng = np.random.default_rng(11)
    n_bits = 16
    src = _rand(1 << n_bits, rng)
    ident = list(range(n_bits))
    pats = {
        "right": list(rng.permutation(10)) + ident[10:],
        "left": ident[:6] + [6 + int(x) for x in rng.permutation(10)],
        "lrl": list(reversed(ident)),
        "identity": ident,
    }
    ctx.reserve((2 << n_bits) + 64)
    ctx.upload(0, src)
    for name, p in pats.items():
        p = [int(x) for x in p]
        ctx.permute(0, (1 << n_bits) + 32, p)
        got = ctx.download((1 << n_bits) + 32, 1 << n_bits)
        assert np.array_equal(got, ko.permute(src, p)), name


def test_permute_involution_full_size(ctx):
    # size-independent property at a BASELINE-sized tensor (2^26 elements,
    # 512 MB): permuting with sigma then sigma^-1 restores the input exactly
    rng = np.random.default_rng(3)
    nbits = 26
    n = 1 << nbits
    src = _rand(n, rng)
    p = [int(x) for x in rng.permutation(nbits)]
    inv = [0] * nbits
    for b, d in enumerate(p):
        inv[d] = b
    ctx.reserve(3 * n)
    ctx.upload(0, src)
    ctx.permute(0, n, p)
    ctx.permute(n, 2 * n, inv)
    assert np.array_equal(ctx.download(2 * n, n), src)


def test_contract_chains_runs_independent_world_lines_in_one_launch():
    """qfx_contract_chains: every chain multiplies small gate tensors one after the
    other (ping-pong buffers), like one grid site of flatten_grid_of_tensors
    (reference src/read_circuit.cpp:548-689); checked step by step against the oracle."""
    from qflex_b200 import _device
    from oracle import kernel_oracle as ko
    rng = np.random.default_rng(2024)
    ctx = _device.Context(0)
    n_chains, steps, max_bits = 5, 9, 12
    buf = 1 << max_bits
    gates_off = 0
    gate_elems = 16
    work = n_chains * steps * gate_elems + 64
    ctx.reserve(work + n_chains * 2 * buf + 64)
    chains, expected = [], []
    for c in range(n_chains):
        ping, pong = work + c * 2 * buf, work + c * 2 * buf + buf
        cur = (rng.uniform(-1, 1, 4) + 1j * rng.uniform(-1, 1, 4)).astype(np.complex64)
        cur_bits = 2
        ctx.upload(ping, cur)
        src, dst = ping, pong
        chain = []
        for s in range(steps):
            # gate [k | n]: contract the running tensor's bit `kp` (and sometimes 2 bits)
            k = 1 if (s % 3) else min(2, cur_bits)
            n = 1 + (s % 2)
            if cur_bits - k + n > max_bits:
                n = 1
            g = (rng.uniform(-1, 1, 1 << (k + n)) + 1j * rng.uniform(-1, 1, 1 << (k + n))).astype(np.complex64)
            goff = gates_off + (c * steps + s) * gate_elems
            ctx.upload(goff, g)
            pos = [int(x) for x in rng.permutation(cur_bits)]
            a_k, a_m = pos[:k], pos[k:]
            m = cur_bits - k
            pc = [int(x) for x in rng.permutation(m + n)]
            t = dict(a_m=a_m, a_k=a_k, b_n=list(range(n)), b_k=list(range(n, n + k)),
                     c_m=pc[:m], c_n=pc[m:])
            chain.append(_device.make_desc(src, goff, dst, t["a_m"], t["c_m"], t["b_n"], t["c_n"],
                                           t["a_k"], t["b_k"]))
            cur = ko.contract(cur, g, **t)
            cur_bits = m + n
            src, dst = dst, src
        chains.append(chain)
        expected.append((src, cur))
    before = ctx.stats()["launches_total"]
    ctx.contract_chains(chains)
    assert ctx.stats()["launches_total"] == before + 1
    for off, want in expected:
        got = ctx.download(off, want.size)
        assert support.rel_err(got, want) < 1e-5
    ctx.close()


def test_copy_and_scalar_accumulate(ctx):
    """qfx_copy (Tensor::operator= / _copy, src/tensor.cpp:107-139) and
    qfx_scalar_accumulate (`amplitudes[i] += *output->data()`,
    src/contraction_utils.cpp:348-352) called directly."""
    rng = np.random.default_rng(5)
    n = 70001  # odd length, not a multiple of the vector width
    a = _rand(n, rng)
    ctx.reserve(3 * n + 64)
    ctx.upload(0, a)
    ctx.upload(n + 7, np.zeros(n, dtype=np.complex64))
    ctx.copy(0, n + 7, n)
    assert np.array_equal(ctx.download(n + 7, n), a)
    assert np.array_equal(ctx.download(0, n), a)
    ctx.copy(3, 2 * n + 20, 1)
    assert ctx.download(2 * n + 20, 1)[0] == a[3]
    ctx.amps_reserve(4)
    for slot, offs in ((0, [1, 5, 9]), (3, [2, 2]), (2, [])):
        for off in offs:
            ctx.scalar_accumulate(off, slot)
    got = ctx.amps(4)
    want = np.array([a[1].astype(np.complex128) + a[5] + a[9], 0, 0, 2 * a[2].astype(np.complex128)])
    assert np.array_equal(got, want)  # complex64 -> complex128 sums of <= 3 terms are exact
    with pytest.raises(Exception):
        ctx.scalar_accumulate(1, 1 << 40)  # slot outside the table
    with pytest.raises(Exception):
        ctx.copy(0, 1 << 40, 8)          # destination outside the arena


def test_headline_shape_full_size_rows_against_oracle(ctx):
    """The class that carries 70 % of bristlecone_70 1-32-1: 2^20 x 256 x 256 on the
    CTA-pair tensor-core kernel, at FULL size; 2048 rows sampled over the whole range
    are recomputed by the oracle in complex128."""
    from qflex_b200 import _device
    m, n, k = 20, 8, 8
    rng = np.random.default_rng(2032)
    t = dict(a_k=list(range(k)), a_m=list(range(k, k + m)), b_n=list(range(n)),
             b_k=list(range(n, n + k)), c_n=list(range(n)), c_m=list(range(n, n + m)))
    sa, sb, sc = 1 << (m + k), 1 << (n + k), 1 << (m + n)
    a = (rng.standard_normal(2 * sa, dtype=np.float32)).view(np.complex64)
    b = (rng.standard_normal(2 * sb, dtype=np.float32)).view(np.complex64)
    oa, ob, oc = 0, sa + 32, sa + sb + 64
    ctx.reserve(oc + sc + 64)
    ctx.upload(oa, a)
    ctx.upload(ob, b)
    before = ctx.stats()
    ctx.contract(_device.make_desc(oa, ob, oc, t["a_m"], t["c_m"], t["b_n"], t["c_n"],
                                   t["a_k"], t["b_k"]))
    after = ctx.stats()
    assert after["launches_gemm_tc"] - before["launches_gemm_tc"] == 1
    rows = np.unique(np.concatenate([rng.integers(0, 1 << m, 2040), [0, 127, 128, 255, 256,
                                                                      (1 << m) - 1, (1 << m) - 256, 1 << 19]]))
    A = a.reshape(1 << m, 1 << k)
    B = b.reshape(1 << k, 1 << n).astype(np.complex128)
    worst = 0.0
    for r in rows:
        got = ctx.download(oc + (int(r) << n), 1 << n)
        want = A[r].astype(np.complex128) @ B
        worst = max(worst, support.rel_err(got, want))
    assert worst < 2e-5, worst


def test_resident_program_of_chains_with_slice_dependent_bases(ctx):
    """qfx_program_create / qfx_program_run: chains of small contractions kept on the
    device, operand bases moved per run by bit fields of the cut values (projected gate
    tensors of the cut-aware flatten) -- against the same steps issued one by one
    through qfx_contract with the bases resolved on the host, and the oracle."""
    import ctypes
    from qflex_b200 import _device
    lib = _device.load()
    rng = np.random.default_rng(77)
    n_chains, steps_per_chain = 5, 4
    # arena: a pool of 8 "gate" tensors of 4x2x2 (16 elements: leading index = projected
    # leg with 4 values), then per chain two ping-pong buffers of 64 elements
    gates = _rand(8 * 16, rng)
    ctx.reserve(8 * 16 + n_chains * 2 * 64 + 64)
    ctx.upload(0, gates)
    seeds = _rand(n_chains * 4, rng)
    descs, chain_first, terms, term_first = [], [0], [], [0]
    plain = []  # (chain, desc fields, var a, var b) for the one-by-one run
    for c in range(n_chains):
        ping = 8 * 16 + c * 128
        pong = ping + 64
        ctx.upload(ping, np.concatenate([seeds[4 * c: 4 * c + 4], np.zeros(60, np.complex64)]))
        src, dst = ping, pong
        for s_ in range(steps_per_chain):
            g = int(rng.integers(0, 8))
            # running tensor [m(1 bit) | k(1 bit)] x gate slice [k | n]: C = [m | n]
            d = _device.make_desc(src, 16 * g, dst, [1], [1], [0], [0], [0], [1])
            cut = int(rng.integers(0, 3))
            frm = int(rng.integers(0, 2))
            va = []
            vb = [(cut, 2, frm, 2)]  # gate base += ((value >> frm) & 3) << 2
            descs.append(d)
            for var in (va, vb):
                terms.extend(var)
                term_first.append(len(terms))
            plain.append((d, va, vb))
            src, dst = dst, src
        chain_first.append(len(descs))
    DescArr = _device.ContractDesc * len(descs)

    class VarTerm(ctypes.Structure):
        _fields_ = [("cut", ctypes.c_int32), ("shift", ctypes.c_int32), ("from_", ctypes.c_int32),
                    ("width", ctypes.c_int32)]
    tarr = (VarTerm * max(1, len(terms)))(*[VarTerm(*t) for t in terms])
    prog = ctypes.c_void_p()
    rc = lib.qfx_program_create(ctx.h, DescArr(*descs), (ctypes.c_int32 * len(chain_first))(*chain_first),
                                n_chains, tarr, (ctypes.c_int32 * len(term_first))(*term_first),
                                ctypes.byref(prog))
    assert rc == 0, lib.qfx_last_error()
    for values in ([0, 0, 0], [3, 1, 2], [5, 7, 6]):
        # reference: one by one with host-resolved bases
        for c in range(n_chains):
            ctx.upload(8 * 16 + c * 128, np.concatenate([seeds[4 * c: 4 * c + 4], np.zeros(124, np.complex64)]))
        for d, va, vb in plain:
            d2 = _device.ContractDesc.from_buffer_copy(d)
            for (cut, shift, frm, width) in vb:
                d2.off_b += ((values[cut] >> frm) & ((1 << width) - 1)) << shift
            ctx.contract(d2)
        want = ctx.download(8 * 16, n_chains * 128)
        for c in range(n_chains):
            ctx.upload(8 * 16 + c * 128, np.concatenate([seeds[4 * c: 4 * c + 4], np.zeros(124, np.complex64)]))
        before = ctx.stats()["launches_total"]
        rc = lib.qfx_program_run(ctx.h, prog, (ctypes.c_uint32 * 3)(*values), 3)
        assert rc == 0, lib.qfx_last_error()
        assert ctx.stats()["launches_total"] - before == 1
        got = ctx.download(8 * 16, n_chains * 128)
        # (another kernel sums the two products in another order: equal to rounding)
        assert support.rel_err(got, want) < 1e-6, values
        assert np.linalg.norm(got) > 0
    # a var term that names a cut the caller did not pass is refused
    assert lib.qfx_program_run(ctx.h, prog, (ctypes.c_uint32 * 1)(0), 1) != 0
    assert lib.qfx_program_destroy(ctx.h, prog) == 0


@pytest.mark.parametrize("m,n,k", [(10, 9, 1), (12, 12, 1), (9, 10, 2), (11, 8, 0), (13, 6, 2), (4, 15, 1)])
def test_contract_small_k_long_m_and_n_output_ordered(ctx, m, n, k):
    """Roots of the tree-ordered world lines: K <= 4, both free sides long, the result
    written in an arbitrary (grid tensor) layout -- enumerated in C-address order."""
    rng = np.random.default_rng(31 * m + 7 * n + k)
    # result layout in groups of bits (bonds), alternately from the two operands
    order = []
    am, bn = list(range(m)), list(range(n))
    while am or bn:
        for src, tag in ((am, "m"), (bn, "n")):
            take = int(rng.integers(1, 5))
            order += [(tag, src.pop(0)) for _ in range(min(take, len(src)))]
    c_m, c_n = [0] * m, [0] * n
    for pos, (tag, i) in enumerate(order):
        (c_m if tag == "m" else c_n)[i] = pos
    pa, pb = rng.permutation(m + k), rng.permutation(n + k)
    t = dict(a_m=sorted(int(x) for x in pa[:m]), a_k=[int(x) for x in pa[m:]],
             b_n=sorted(int(x) for x in pb[:n]), b_k=[int(x) for x in pb[n:]], c_m=c_m, c_n=c_n)
    assert _run_contract(ctx, t, rng) < TOL
    assert _run_contract(ctx, _random_tables(rng, m, n, k), rng) < TOL


@pytest.mark.parametrize("dims,order", [([3, 2, 5], [2, 0, 1]), ([2, 4, 3, 1], [3, 1, 0, 2]),
                                        ([6, 7], [1, 0]), ([5, 1, 2, 3, 2], [4, 2, 0, 3, 1]),
                                        ([3, 3, 3, 3], [0, 1, 2, 3]), ([12, 10, 9], [1, 2, 0]),
                                        ([7, 33, 5, 64, 3], [3, 0, 4, 1, 2]), ([1], [0]),
                                        ([2] * 12, [5, 0, 11, 3, 7, 1, 9, 2, 10, 4, 8, 6])])
def test_reorder_general_dimensions(ctx, dims, order):
    """qfx_reorder_general (Tensor::_naive_reorder, src/tensor.cpp:377-492): bit-exact
    against the oracle's transpose (itself pinned to the reference in tests/test_oracle.py)."""
    rng = np.random.default_rng(sum(dims) + len(dims))
    total = int(np.prod(dims))
    data = _rand(total, rng)
    idx = ["a%d" % i for i in range(len(dims))]
    want, _ = ko.reorder_named(data, idx, dims, [idx[p] for p in order])
    ctx.reserve(2 * total + 64)
    ctx.upload(3, data)
    ctx.upload(total + 16, np.zeros(total, dtype=np.complex64))
    ctx.reorder_general(3, total + 16, dims, order)
    assert np.array_equal(ctx.download(total + 16, total), want)
    assert np.array_equal(ctx.download(3, total), data)


def test_reorder_general_rejects_bad_arguments(ctx):
    ctx.reserve(1 << 12)
    with pytest.raises(Exception, match="cannot be zero"):
        ctx.reorder_general(0, 100, [], [])
    with pytest.raises(Exception, match="permutation"):
        ctx.reorder_general(0, 100, [3, 2], [0, 0])
    with pytest.raises(Exception, match="overlap"):
        ctx.reorder_general(0, 5, [3, 2], [1, 0])
    with pytest.raises(Exception, match="outside the arena"):
        ctx.reorder_general(0, 1 << 39, [1 << 20, 1 << 19], [1, 0])  # (a parked arena may be larger than asked)
    with pytest.raises(Exception, match="bad dimension"):
        ctx.reorder_general(0, 100, [3, 0], [1, 0])
