"""TEST INFRASTRUCTURE: ctypes access to the UNMODIFIED reference built by
oracle/Makefile into oracle/_ref (libqflex_ref.so = the reference's translation units +
oracle/ref_capi.cpp).  Used by tests/ as the checker and by bench.py's cpu_baseline /
--impl reference legs as the timed CPU arm; never by the product (qflex_b200/)."""
import ctypes
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libqflex_ref.so")

_ref = None


def reference_available():
    return os.path.exists(REF_LIB)


def reference_library():
    global _ref
    if _ref is None:
        lib = ctypes.CDLL(REF_LIB)
        lib.ref_evaluate.restype = ctypes.c_long
        lib.ref_multiply.restype = ctypes.c_long
        lib.ref_reorder.restype = ctypes.c_long
        _ref = lib
    return _ref


def reference_set_threads(n):
    reference_library().ref_set_blas_threads(int(n))


def reference_evaluate(circuit, ordering, grid, initial="", final="", memory=64 << 30,
                       cap=1 << 16):
    """The reference's own EvaluateCircuit (oracle/ref_capi.cpp)."""
    lib = reference_library()
    amps = np.zeros(2 * cap)
    states = ctypes.create_string_buffer(cap * 96)
    err = ctypes.create_string_buffer(1 << 16)
    secs = ctypes.c_double()
    n = lib.ref_evaluate(circuit.encode(), ordering.encode(), grid.encode(), initial.encode(),
                         final.encode(), ctypes.c_ulong(memory), states,
                         ctypes.c_ulong(len(states)), amps.ctypes.data_as(ctypes.c_void_p),
                         ctypes.c_ulong(cap), ctypes.byref(secs), err, ctypes.c_ulong(len(err)))
    if n < 0:
        raise RuntimeError(err.value.decode())
    vals = amps[: 2 * n].view(np.complex128)
    return list(zip(states.value.decode().split("\n"), [complex(v) for v in vals])), secs.value


def reference_multiply(a, a_idx, a_dims, b, b_idx, b_dims):
    """The reference's own qflex::multiply on caller data."""
    lib = reference_library()
    a = np.ascontiguousarray(a, dtype=np.complex64).copy()
    b = np.ascontiguousarray(b, dtype=np.complex64).copy()
    ad = (ctypes.c_ulong * len(a_dims))(*a_dims)
    bd = (ctypes.c_ulong * len(b_dims))(*b_dims)
    cap = int(np.prod(a_dims, dtype=np.int64) * np.prod(b_dims, dtype=np.int64))
    c = np.zeros(max(cap, 1), dtype=np.complex64)
    names = ctypes.create_string_buffer(4096)
    cd = (ctypes.c_ulong * 64)()
    err = ctypes.create_string_buffer(1 << 14)
    n = lib.ref_multiply(";".join(a_idx).encode(), ad, a.ctypes.data_as(ctypes.c_void_p),
                         ";".join(b_idx).encode(), bd, b.ctypes.data_as(ctypes.c_void_p),
                         c.ctypes.data_as(ctypes.c_void_p), ctypes.c_ulong(c.size), names,
                         ctypes.c_ulong(len(names)), cd, err, ctypes.c_ulong(len(err)))
    if n < 0:
        raise RuntimeError(err.value.decode())
    dims = [int(cd[i]) for i in range(n)]
    idx = names.value.decode().split(";") if n else []
    return c[: int(np.prod(dims, dtype=np.int64)) if dims else 1], idx, dims


def reference_reorder(data, idx, dims, new_idx):
    lib = reference_library()
    d = np.ascontiguousarray(data, dtype=np.complex64).copy()
    dd = (ctypes.c_ulong * len(dims))(*dims)
    err = ctypes.create_string_buffer(1 << 14)
    rc = lib.ref_reorder(";".join(idx).encode(), dd, d.ctypes.data_as(ctypes.c_void_p),
                         ";".join(new_idx).encode(), err, ctypes.c_ulong(len(err)))
    if rc < 0:
        raise RuntimeError(err.value.decode())
    return d
