"""ctypes binding of libqflex_b200.so (include/qflex_b200.h): the device C ABI.

Used by the kernel parity tests and bench.py's kernel micro-benchmarks.  There
is no CPU path: loading fails loudly when the library is missing and
``Context()`` fails loudly when no sm_100 GPU is present.
"""
import ctypes
import os

import numpy as np

MAX_BITS = 40
_HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_DEVICE_LIB = os.path.join(_HERE, "libqflex_b200.so")

EXPORTS = (
    "qfx_last_error", "qfx_device_count", "qfx_create", "qfx_destroy", "qfx_sync",
    "qfx_arena_reserve", "qfx_arena_size", "qfx_upload", "qfx_download", "qfx_copy",
    "qfx_contract", "qfx_dot_accumulate", "qfx_scalar_accumulate", "qfx_tensor_accumulate",
    "qfx_amps_reserve",
    "qfx_amps_download", "qfx_permute", "qfx_timer_start", "qfx_timer_stop_ms",
    "qfx_stats_get", "qfx_stats_reset", "qfx_profile_begin", "qfx_profile_end",
)


class ContractDesc(ctypes.Structure):
    _fields_ = [
        ("off_a", ctypes.c_uint64), ("off_b", ctypes.c_uint64), ("off_c", ctypes.c_uint64),
        ("m_bits", ctypes.c_int32), ("n_bits", ctypes.c_int32), ("k_bits", ctypes.c_int32),
        ("a_m", ctypes.c_int8 * MAX_BITS), ("c_m", ctypes.c_int8 * MAX_BITS),
        ("b_n", ctypes.c_int8 * MAX_BITS), ("c_n", ctypes.c_int8 * MAX_BITS),
        ("a_k", ctypes.c_int8 * MAX_BITS), ("b_k", ctypes.c_int8 * MAX_BITS),
        ("precision", ctypes.c_int32),
    ]


class Stats(ctypes.Structure):
    _fields_ = [
        ("launches_total", ctypes.c_uint64), ("launches_gemm_ffma", ctypes.c_uint64),
        ("launches_gemm_tc", ctypes.c_uint64), ("launches_reduce", ctypes.c_uint64),
        ("launches_permute", ctypes.c_uint64), ("launches_copy", ctypes.c_uint64),
        ("flops", ctypes.c_double), ("bytes", ctypes.c_double),
        ("launches_gemm_tc_fused", ctypes.c_uint64),
        ("launches_gemm_tc_pair", ctypes.c_uint64),
    ]


class DeviceError(RuntimeError):
    pass


def load(path=DEFAULT_DEVICE_LIB):
    # A/B experiments (tools/ only): alternative build of the same library
    path = os.environ.get("QFLEX_B200_DEVICE_LIB", path)
    if not os.path.exists(path):
        raise DeviceError("%s not found: run __graft_entry__.build() (no CPU fallback)" % path)
    c = ctypes.CDLL(path)
    c.qfx_last_error.restype = ctypes.c_char_p
    u64, vp, i32 = ctypes.c_uint64, ctypes.c_void_p, ctypes.c_int32
    c.qfx_create.argtypes = [ctypes.c_int, ctypes.POINTER(vp)]
    c.qfx_destroy.argtypes = [vp]
    c.qfx_sync.argtypes = [vp]
    c.qfx_arena_reserve.argtypes = [vp, u64]
    c.qfx_upload.argtypes = [vp, u64, vp, u64]
    c.qfx_download.argtypes = [vp, u64, vp, u64]
    c.qfx_copy.argtypes = [vp, u64, u64, u64]
    c.qfx_contract.argtypes = [vp, ctypes.POINTER(ContractDesc)]
    c.qfx_contract_pair_static.argtypes = [vp, ctypes.POINTER(ContractDesc)]
    c.qfx_dot_accumulate.argtypes = [vp, ctypes.POINTER(ContractDesc), u64]
    c.qfx_contract_chains.argtypes = [vp, ctypes.POINTER(ContractDesc),
                                      ctypes.POINTER(ctypes.c_int32), i32]
    c.qfx_program_create.argtypes = [vp, vp, vp, i32, vp, vp, ctypes.POINTER(vp)]
    c.qfx_program_run.argtypes = [vp, vp, vp, i32]
    c.qfx_program_destroy.argtypes = [vp, vp]
    c.qfx_trim_cache.argtypes = [ctypes.c_int]
    c.qfx_bind_thread.argtypes = [vp]
    c.qfx_scalar_accumulate.argtypes = [vp, u64, u64]
    c.qfx_tensor_accumulate.argtypes = [vp, u64, i32, ctypes.POINTER(ctypes.c_uint32), u64]
    c.qfx_amps_reserve.argtypes = [vp, u64]
    c.qfx_amps_download.argtypes = [vp, vp, u64]
    c.qfx_permute.argtypes = [vp, u64, u64, i32, vp]
    c.qfx_reorder_general.argtypes = [vp, u64, u64, i32, vp, vp]
    c.qfx_timer_start.argtypes = [vp]
    c.qfx_timer_stop_ms.argtypes = [vp, ctypes.POINTER(ctypes.c_float)]
    c.qfx_stats_get.argtypes = [vp, ctypes.POINTER(Stats)]
    c.qfx_stats_reset.argtypes = [vp]
    c.qfx_device_count.argtypes = [ctypes.POINTER(ctypes.c_int)]
    return c


def make_desc(off_a, off_b, off_c, a_m, c_m, b_n, c_n, a_k, b_k, precision=0):
    d = ContractDesc()
    d.off_a, d.off_b, d.off_c = off_a, off_b, off_c
    d.m_bits, d.n_bits, d.k_bits = len(a_m), len(b_n), len(a_k)
    for name, vals in (("a_m", a_m), ("c_m", c_m), ("b_n", b_n), ("c_n", c_n),
                       ("a_k", a_k), ("b_k", b_k)):
        arr = getattr(d, name)
        for i, v in enumerate(vals):
            arr[i] = v
    d.precision = precision
    return d


class Context:
    """One GPU context (arena + amplitude table)."""

    def __init__(self, device=0, lib=None):
        self.c = lib or load()
        h = ctypes.c_void_p()
        self._ck(self.c.qfx_create(device, ctypes.byref(h)))
        self.h = h

    def _ck(self, rc):
        if rc != 0:
            raise DeviceError(self.c.qfx_last_error().decode())

    def reserve(self, n_elems):
        self._ck(self.c.qfx_arena_reserve(self.h, n_elems))

    def upload(self, off, arr):
        a = np.ascontiguousarray(arr, dtype=np.complex64)
        self._ck(self.c.qfx_upload(self.h, off, a.ctypes.data_as(ctypes.c_void_p), a.size))

    def download(self, off, n):
        out = np.empty(n, dtype=np.complex64)
        self._ck(self.c.qfx_download(self.h, off, out.ctypes.data_as(ctypes.c_void_p), n))
        return out

    def contract_chains(self, chains):
        """chains: list of lists of ContractDesc; one launch (qfx_contract_chains)."""
        flat = [d for ch in chains for d in ch]
        arr = (ContractDesc * max(1, len(flat)))(*flat)
        first = [0]
        for ch in chains:
            first.append(first[-1] + len(ch))
        farr = (ctypes.c_int32 * len(first))(*first)
        self._ck(self.c.qfx_contract_chains(self.h, arr, farr, len(chains)))

    def contract(self, desc):
        self._ck(self.c.qfx_contract(self.h, ctypes.byref(desc)))

    def contract_pair_static(self, desc):
        """qfx_contract_pair_static: the CTA-pair kernel with its scratch in a static
        device array (no device allocation inside the call)."""
        self._ck(self.c.qfx_contract_pair_static(self.h, ctypes.byref(desc)))

    def dot_accumulate(self, desc, slot):
        self._ck(self.c.qfx_dot_accumulate(self.h, ctypes.byref(desc), slot))

    def scalar_accumulate(self, off, slot):
        self._ck(self.c.qfx_scalar_accumulate(self.h, off, slot))

    def tensor_accumulate(self, off, slot_stride, slot=0):
        st = (ctypes.c_uint32 * max(1, len(slot_stride)))(*slot_stride)
        self._ck(self.c.qfx_tensor_accumulate(self.h, off, len(slot_stride), st, slot))

    def reorder_general(self, src, dst, dims, new_order):
        """qfx_reorder_general: axes of arbitrary dimensions (Tensor::_naive_reorder)."""
        d = (ctypes.c_uint32 * max(1, len(dims)))(*dims)
        o = (ctypes.c_int32 * max(1, len(new_order)))(*new_order)
        self._ck(self.c.qfx_reorder_general(self.h, src, dst, len(dims), d, o))

    def permute(self, src, dst, dst_pos):
        p = (ctypes.c_int8 * max(1, len(dst_pos)))(*dst_pos)
        self._ck(self.c.qfx_permute(self.h, src, dst, len(dst_pos), p))

    def copy(self, src, dst, n):
        self._ck(self.c.qfx_copy(self.h, src, dst, n))

    def amps_reserve(self, n):
        self._ck(self.c.qfx_amps_reserve(self.h, n))

    def amps(self, n):
        out = np.empty(n, dtype=np.complex128)
        self._ck(self.c.qfx_amps_download(self.h, out.ctypes.data_as(ctypes.c_void_p), n))
        return out

    def sync(self):
        self._ck(self.c.qfx_sync(self.h))

    def timer_start(self):
        self._ck(self.c.qfx_timer_start(self.h))

    def timer_stop_ms(self):
        ms = ctypes.c_float()
        self._ck(self.c.qfx_timer_stop_ms(self.h, ctypes.byref(ms)))
        return ms.value

    def stats(self):
        s = Stats()
        self._ck(self.c.qfx_stats_get(self.h, ctypes.byref(s)))
        return {f: getattr(s, f) for f, _ in Stats._fields_}

    def stats_reset(self):
        self._ck(self.c.qfx_stats_reset(self.h))

    def close(self):
        if self.h:
            self.c.qfx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
