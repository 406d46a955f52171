"""qfx_contract_pair_static: the CTA-pair tensor-core contraction with its scratch in a
static device array (no cudaMalloc inside the call), checked against qfx_contract on
the same operands.  Collected last on purpose."""
import ctypes

import numpy as np
import pytest

from tests import support

pytestmark = pytest.mark.gpu

TOL = 1e-5  # as tests/test_gpu_kernels.py


def _tables(m, n, k):
    # the layout an expand step produces: A = [m | k], B = [k | n], C = [m | n]
    return dict(a_k=list(range(k)), a_m=list(range(k, k + m)), b_n=list(range(n)),
                b_k=list(range(n, n + k)), c_n=list(range(n)), c_m=list(range(n, n + m)))


def _free_bytes(lib):
    free, total = ctypes.c_uint64(0), ctypes.c_uint64(0)
    assert lib.qfx_device_memory(0, ctypes.byref(free), ctypes.byref(total)) == 0
    return free.value


def test_static_scratch_entry_matches_qfx_contract_twice_in_a_row():
    from qflex_b200 import _device
    ctx = _device.Context(0)
    try:
        rng = np.random.default_rng(4242)
        shapes = [(16, 8, 8), (17, 6, 4)]  # B' resident / dual-stage form of the kernel
        plan, top = [], 0
        for m, n, k in shapes:
            sa, sb, sc = 1 << (m + k), 1 << (n + k), 1 << (m + n)
            oa, ob, oc, od = top, top + sa + 32, top + sa + sb + 64, top + sa + sb + sc + 96
            top = od + sc + 32
            plan.append((m, n, k, oa, ob, oc, od))
        ctx.reserve(top)
        want = []
        for m, n, k, oa, ob, oc, od in plan:
            a = rng.standard_normal(2 << (m + k), dtype=np.float32).view(np.complex64)
            b = rng.standard_normal(2 << (n + k), dtype=np.float32).view(np.complex64)
            ctx.upload(oa, a)
            ctx.upload(ob, b)
            junk = np.full(1 << (m + n), 7 + 7j, dtype=np.complex64)
            ctx.upload(oc, junk)
            ctx.upload(od, junk)
            t = _tables(m, n, k)
            before = ctx.stats()["launches_gemm_tc_pair"]
            ctx.contract(_device.make_desc(oa, ob, oc, t["a_m"], t["c_m"], t["b_n"], t["c_n"],
                                           t["a_k"], t["b_k"]))
            assert ctx.stats()["launches_gemm_tc_pair"] == before + 1
            want.append(ctx.download(oc, 1 << (m + n)))
        # the existing entry point against the oracle's arithmetic on a few rows
        m, n, k, oa, ob, oc, od = plan[0]
        A = ctx.download(oa, 1 << (m + k)).reshape(1 << m, 1 << k)
        B = ctx.download(ob, 1 << (n + k)).reshape(1 << k, 1 << n).astype(np.complex128)
        for r in (0, 255, 256, (1 << m) - 1):
            row = want[0][r << n:(r + 1) << n]
            assert support.rel_err(row, A[r].astype(np.complex128) @ B) < 2e-5
        # two calls in a row through the new entry point, different shapes and data: the
        # second call reuses the array the first one filled
        ctx.sync()
        for rep in range(2):
            free_before = _free_bytes(ctx.c)
            for (m, n, k, oa, ob, oc, od), ref in zip(plan, want):
                if rep:
                    ctx.upload(od, np.zeros(1 << (m + n), dtype=np.complex64))
                t = _tables(m, n, k)
                ctx.contract_pair_static(_device.make_desc(oa, ob, od, t["a_m"], t["c_m"],
                                                           t["b_n"], t["c_n"], t["a_k"],
                                                           t["b_k"]))
                got = ctx.download(od, 1 << (m + n))
                assert support.rel_err(got, ref) < TOL
            if rep:  # (the first round may load the kernels' module lazily)
                assert _free_bytes(ctx.c) >= free_before
    finally:
        ctx.close()


def test_static_scratch_entry_rejects_what_does_not_fit():
    from qflex_b200 import _device
    ctx = _device.Context(0)
    try:
        ctx.reserve((1 << 26) + (1 << 14))

        def call(m, n, k):
            t = _tables(m, n, k)
            sa, sb = 1 << (m + k), 1 << (n + k)
            ctx.contract_pair_static(_device.make_desc(0, sa, sa + sb, t["a_m"], t["c_m"],
                                                       t["b_n"], t["c_n"], t["a_k"], t["b_k"]))

        with pytest.raises(Exception, match="exceeds the static scratch"):
            call(12, 9, 8)   # smaller operand 2^17 elements
        with pytest.raises(Exception, match="exceeds the static scratch"):
            call(9, 14, 4)   # roles exchanged: n = 9 after the swap
        with pytest.raises(Exception, match="CTA-pair kernel's range"):
            call(10, 6, 6)   # too few rows for the pairs
        with pytest.raises(Exception, match="overlaps"):
            t = _tables(16, 8, 8)
            ctx.contract_pair_static(_device.make_desc(0, 1 << 24, 8, t["a_m"], t["c_m"],
                                                       t["b_n"], t["c_n"], t["a_k"], t["b_k"]))
    finally:
        ctx.close()
