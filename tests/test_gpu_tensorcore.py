"""Parity of the tcgen05 split-bf16 contraction kernels (fused in-kernel
conversion, tc_fused.cu; pre-split fallback, tc_gemm.cu) with the oracle.

Tolerance: the bf16 hi+mid split keeps 16-17 mantissa bits per factor, so one
contraction is within ~4e-6 normwise of the complex128 oracle (measured,
tools/tc_error.py); the test bound is 2e-5.  The end-to-end bound (1e-4 on the
amplitude vector, BASELINE north_star) is asserted in test_gpu_e2e.py."""
import numpy as np
import pytest

from oracle import kernel_oracle as ko
from tests import support

pytestmark = pytest.mark.gpu

TOL_TC = 2e-5


def _rand(n, rng):
    return (rng.uniform(-1, 1, n) + 1j * rng.uniform(-1, 1, n)).astype(np.complex64)


@pytest.fixture(scope="module")
def ctx():
    from qflex_b200 import _device
    c = _device.Context(0)
    yield c
    c.close()


def _run(ctx, t, rng, jitter=0, precision=0):
    from qflex_b200 import _device
    m, n, k = len(t["a_m"]), len(t["b_n"]), len(t["a_k"])
    sa, sb, sc = 1 << (m + k), 1 << (n + k), 1 << (m + n)
    a, b = _rand(sa, rng), _rand(sb, rng)
    oa = jitter
    ob = oa + sa + 32 + jitter
    oc = ob + sb + 32 + jitter
    ctx.reserve(oc + sc + 64)
    ctx.upload(oa, a)
    ctx.upload(ob, b)
    ctx.upload(oc, np.zeros(sc, dtype=np.complex64))
    before = ctx.stats()
    ctx.contract(_device.make_desc(oa, ob, oc, t["a_m"], t["c_m"], t["b_n"], t["c_n"],
                                   t["a_k"], t["b_k"], precision=precision))
    got = ctx.download(oc, sc)
    after = ctx.stats()
    want = ko.contract(a, b, acc_dtype=np.complex128, **t)
    used = {key: after[key] - before[key] for key in
            ("launches_gemm_tc", "launches_gemm_tc_fused", "launches_gemm_ffma")}
    return support.rel_err(got, want), used


def _natural(m, n, k):
    """A = [m | k], B = [k | n], C = [m | n] (k / n lowest)."""
    return dict(a_k=list(range(k)), a_m=list(range(k, k + m)),
                b_n=list(range(n)), b_k=list(range(n, n + k)),
                c_n=list(range(n)), c_m=list(range(n, n + m)))


def _grouped(m, n, k, k_groups):
    """Bristlecone-style patch layout: A's k bits sit in 4-bit groups at the
    address groups `k_groups`; m bits fill the rest in order (the layouts of
    bristlecone_70 1-32-1, e.g. a_k = 4..7, 24..27)."""
    a_k = [4 * g + i for g in k_groups for i in range(4)][:k]
    assert len(a_k) == k
    a_m = [p for p in range(m + k) if p not in a_k]
    t = _natural(m, n, k)
    t["a_k"], t["a_m"] = a_k, a_m
    return t


@pytest.mark.parametrize("m,n,k", [(14, 8, 8), (12, 8, 8), (16, 4, 4), (15, 5, 5), (13, 6, 6),
                                   (11, 7, 9), (14, 4, 8), (13, 8, 4), (10, 8, 12)])
def test_fused_natural_layout(ctx, m, n, k):
    rng = np.random.default_rng(100 * m + 10 * n + k)
    err, used = _run(ctx, _natural(m, n, k), rng)
    assert err < TOL_TC
    assert used["launches_gemm_tc_fused"] == 1, used


@pytest.mark.parametrize("m,n,k,groups", [
    (14, 8, 8, (1, 4)),   # a_k = 4..7, 16..19  (expand A (7,5) family)
    (14, 8, 8, (0, 4)),   # a_k = 0..3, 16..19
    (14, 8, 8, (2, 3)),   # a_k = 8..15, m bits below and above
    (16, 4, 4, (1,)),     # a_k = 4..7       (2^24 x 16 x 16 family)
    (16, 4, 4, (4,)),     # a_k = 16..19: tile rows are A's lowest 7 bits
    (16, 4, 4, (0,)),
    (12, 8, 8, (0, 1)),
])
def test_fused_grouped_layouts_of_the_bristlecone_plans(ctx, m, n, k, groups):
    rng = np.random.default_rng(hash((m, n, k, groups)) % (1 << 31))
    err, used = _run(ctx, _grouped(m, n, k, groups), rng)
    assert err < TOL_TC
    assert used["launches_gemm_tc_fused"] == 1, used


def test_fused_scattered_c(ctx):
    # C addressed through tables (merge results are not always [m | n])
    rng = np.random.default_rng(77)
    t = _grouped(15, 6, 4, (1,))
    pc = [int(x) for x in rng.permutation(15 + 6)]
    t["c_m"], t["c_n"] = pc[:15], pc[15:]
    err, used = _run(ctx, t, rng)
    assert err < TOL_TC
    assert used["launches_gemm_tc_fused"] == 1, used


@pytest.mark.parametrize("m,n,k", [(12, 7, 7), (13, 5, 6), (11, 8, 8)])
def test_random_bit_placement_any_tensor_core_path(ctx, m, n, k):
    # arbitrary placements: fused when the A tile is expressible as a TMA box,
    # otherwise the pre-split tensor-core path; both must hold the tolerance
    rng = np.random.default_rng(31 * m + 7 * n + k)
    pa, pb, pc = rng.permutation(m + k), rng.permutation(n + k), rng.permutation(m + n)
    t = dict(a_m=[int(x) for x in pa[:m]], a_k=[int(x) for x in pa[m:]],
             b_k=[int(x) for x in pb[:k]], b_n=[int(x) for x in pb[k:]],
             c_m=[int(x) for x in pc[:m]], c_n=[int(x) for x in pc[m:]])
    err, used = _run(ctx, t, rng)
    assert err < TOL_TC
    assert used["launches_gemm_tc"] + used["launches_gemm_ffma"] == 1, used


def test_odd_base_offsets_fall_back(ctx):
    # cut projections can shift an operand base to an odd element: TMA needs
    # 16-byte alignment, so the fused kernel must decline and a fallback run
    rng = np.random.default_rng(9)
    err, used = _run(ctx, _natural(13, 6, 6), rng, jitter=1)
    assert err < TOL_TC
    assert used["launches_gemm_tc_fused"] == 0, used


def test_swapped_roles(ctx):
    # n_bits > m_bits: the launcher swaps operands so the larger side streams
    rng = np.random.default_rng(4)
    t = _natural(6, 13, 6)
    err, used = _run(ctx, t, rng)
    assert err < TOL_TC
    assert used["launches_gemm_tc"] == 1, used


def test_ffma_precision_request_is_honoured(ctx):
    rng = np.random.default_rng(5)
    err, used = _run(ctx, _natural(13, 6, 6), rng, precision=1)
    assert err < 1e-5
    assert used["launches_gemm_tc"] == 0 and used["launches_gemm_ffma"] == 1, used


# ---- CTA-pair kernel (tc_pair.cu): k, n <= 8 bits and at least 74 256-row tiles
@pytest.mark.parametrize("m,n,k,groups", [
    (15, 8, 8, None),      # two n tiles, 8 k blocks: the 2^20 x 256 x 256 family
    (15, 8, 8, (1, 4)),    # a_k = 4..7, 16..19
    (15, 8, 8, (0, 2)),
    (16, 7, 8, None),      # one n tile of 256 columns, two accumulator sets
    (16, 6, 7, (1, 3)),
    (17, 4, 4, (1,)),      # 2^24 x 16 x 16 family: resident B', half-filled k block
    (17, 4, 4, (4,)),
    (16, 5, 5, None),
    (16, 8, 4, None),
    (15, 4, 8, (0, 1)),
])
def test_pair_kernel_matches_oracle(ctx, m, n, k, groups):
    rng = np.random.default_rng(1000 * m + 10 * n + k)
    t = _natural(m, n, k) if groups is None else _grouped(m, n, k, groups)
    err, used = _run(ctx, t, rng)
    assert err < TOL_TC
    assert ctx.stats()["launches_gemm_tc_pair"] >= 1


def test_pair_kernel_scattered_c_and_repeatability(ctx):
    rng = np.random.default_rng(78)
    t = _grouped(15, 6, 6, (1, 2))
    pc = [int(x) for x in rng.permutation(15 + 6)]
    t["c_m"], t["c_n"] = pc[:15], pc[15:]
    before = ctx.stats()["launches_gemm_tc_pair"]
    e1, _ = _run(ctx, t, np.random.default_rng(5))
    e2, _ = _run(ctx, t, np.random.default_rng(5))
    assert e1 < TOL_TC and e1 == e2
    assert ctx.stats()["launches_gemm_tc_pair"] == before + 2


@pytest.mark.parametrize("m,n,k,groups,c_layout", [
    (17, 5, 4, None, "natural"), (17, 6, 4, (1,), "natural"), (18, 4, 4, (2,), "natural"),
    (17, 4, 4, (1,), "scattered"), (17, 6, 4, None, "scattered"), (17, 5, 4, (3,), "n_high_m_low"),
])
def test_pair_kernel_dual_stage_epilogue(ctx, m, n, k, groups, c_layout):
    """K = 16 (dual stages): epilogue warps 0-3 / 4-7 drain the stage's two accumulators,
    one whole C row per thread (16, 32 and 64 complex columns), natural and scattered C."""
    rng = np.random.default_rng(52000 + 100 * m + 10 * n + k + len(c_layout))
    t = _natural(m, n, k) if groups is None else _grouped(m, n, k, groups)
    if c_layout == "scattered":
        pc = [int(x) for x in rng.permutation(m + n)]
        t["c_m"], t["c_n"] = pc[:m], pc[m:]
    elif c_layout == "n_high_m_low":  # n bits above bit 5: C address bits 0, 1 are not lane rows
        t["c_n"] = list(range(m, m + n))
        t["c_m"] = [int(x) for x in rng.permutation(m)]
    before = ctx.stats()["launches_gemm_tc_pair"]
    err, used = _run(ctx, t, rng)
    assert err < TOL_TC
    assert ctx.stats()["launches_gemm_tc_pair"] == before + 1


# ---- CTA-pair kernel for long K / wide N (tc_pairlong.cu): k > 8 bits or n > 8 bits and
# at least two waves of 256-row work items
@pytest.mark.parametrize("m,n,k,groups,c_layout", [
    (14, 9, 9, None, "natural"),        # 4 n tiles, 16 k blocks (4 chunks)
    (15, 8, 10, (0, 2, 4), "natural"),  # 2 n tiles, 32 k blocks, grouped A layout
    (16, 6, 9, None, "natural"),        # one n tile of 128 columns: 4 accumulator stages
    (14, 12, 5, None, "natural"),       # wide N, one k block
    (15, 9, 6, (1, 3), "scattered"),    # k blocks = 2 < chunk
    (14, 10, 9, None, "transposed"),
    (16, 9, 11, None, "natural"),       # 64 k blocks: the bias compensation of 16 chunks
])
def test_pair_kernel_long_k_wide_n(ctx, m, n, k, groups, c_layout):
    rng = np.random.default_rng(61000 + 100 * m + 10 * n + k)
    t = _natural(m, n, k) if groups is None else _grouped(m, n, k, groups)
    if c_layout == "scattered":
        pc = [int(x) for x in rng.permutation(m + n)]
        t["c_m"], t["c_n"] = pc[:m], pc[m:]
    elif c_layout == "transposed":
        t["c_m"], t["c_n"] = list(range(m)), list(range(m, m + n))
    before = ctx.stats()["launches_gemm_tc_pair"]
    err, used = _run(ctx, t, rng)
    assert err < TOL_TC
    assert ctx.stats()["launches_gemm_tc_pair"] == before + 1, used


@pytest.mark.parametrize("m,n,k,c_layout", [(12, 8, 14, "natural"), (11, 8, 15, "scattered"),
                                            (12, 7, 14, "transposed")])
def test_pair_kernel_long_k_split_k(ctx, m, n, k, c_layout):
    """Few 256-row work items and a long K (the headline's 2^12 x 256 x 2^16 merges): K is
    split over the pair kernel's work items, slabs summed by tc_splitk_finish."""
    rng = np.random.default_rng(62000 + 100 * m + 10 * n + k)
    t = _natural(m, n, k)
    if c_layout == "scattered":
        pc = [int(x) for x in rng.permutation(m + n)]
        t["c_m"], t["c_n"] = pc[:m], pc[m:]
    elif c_layout == "transposed":
        t["c_m"], t["c_n"] = list(range(m)), list(range(m, m + n))
    before = ctx.stats()
    err, used = _run(ctx, t, rng)
    after = ctx.stats()
    assert err < TOL_TC
    assert after["launches_gemm_tc_pair"] == before["launches_gemm_tc_pair"] + 1, used
    assert after["launches_reduce"] == before["launches_reduce"] + 1  # the slab sum


# ---- split-K in tc_fused.cu: few output tiles, long K (the final merges)
@pytest.mark.parametrize("m,n,k,scatter", [(10, 8, 12, False), (9, 6, 13, True), (8, 8, 14, False),
                                           (12, 8, 12, True)])
def test_fused_split_k(ctx, m, n, k, scatter):
    rng = np.random.default_rng(7000 + 100 * m + 10 * n + k)
    t = _natural(m, n, k)
    if scatter:
        pc = [int(x) for x in rng.permutation(m + n)]
        t["c_m"], t["c_n"] = pc[:m], pc[m:]
    before = ctx.stats()
    e1, used = _run(ctx, t, np.random.default_rng(11))
    e2, _ = _run(ctx, t, np.random.default_rng(11))
    after = ctx.stats()
    assert e1 < TOL_TC and e1 == e2          # fixed-order slab sum: deterministic
    assert used["launches_gemm_tc_fused"] == 1, used
    assert after["launches_reduce"] - before["launches_reduce"] == 2  # one finish per run


# ---- transposed C (rows of the tile are C's fast dimension: what a role swap gives)
@pytest.mark.parametrize("m,n,k,groups", [
    (12, 8, 8, None), (13, 6, 9, None), (14, 8, 8, (1, 4)), (13, 5, 8, None),
    (15, 8, 8, None), (16, 7, 8, (0, 2)), (17, 5, 5, None), (17, 4, 4, (1,)),
])
def test_transposed_c_layout(ctx, m, n, k, groups):
    rng = np.random.default_rng(31000 + 100 * m + 10 * n + k)
    t = _natural(m, n, k) if groups is None else _grouped(m, n, k, groups)
    t["c_m"], t["c_n"] = list(range(m)), list(range(m, m + n))
    err, used = _run(ctx, t, rng)
    assert err < TOL_TC
    assert used["launches_gemm_tc_fused"] == 1, used


def test_role_swap_produces_transposed_c_and_matches(ctx):
    # n_bits > m_bits with natural C: the launcher swaps the operands, C becomes transposed
    rng = np.random.default_rng(41)
    err, used = _run(ctx, _natural(8, 14, 8), rng)
    assert err < TOL_TC
    assert used["launches_gemm_tc"] == 1, used


@pytest.mark.parametrize("m,n,k", [(16, 7, 7), (15, 7, 8), (13, 7, 7), (12, 8, 7)])
def test_b_prime_too_large_to_stay_resident_is_streamed(ctx, m, n, k):
    # N' = 256 / K' = 256..512: all of B' would need > 227 KB of shared memory; the
    # in-kernel-conversion kernels must stream it, not hand over to the pre-split path
    rng = np.random.default_rng(52000 + 100 * m + 10 * n + k)
    err, used = _run(ctx, _natural(m, n, k), rng)
    assert err < TOL_TC
    assert used["launches_gemm_tc_fused"] == 1, used
