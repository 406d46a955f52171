"""Systematic part of the tensor-core path's error: the mean relative PROJECTION of
(got - exact) on exact, per contraction, for the shapes of bristlecone_70 1-32-1.
A round-toward-zero accumulator shrinks every output by the same expected factor, and
such a bias adds up linearly along a chain of contractions (the random part only in
quadrature).  Prints bias and normwise error for tensor-core and FFMA paths."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from qflex_b200 import _device  # noqa: E402


def main():
    ctx = _device.Context(0)
    rng = np.random.default_rng(0)
    shapes = [(16, 8, 8), (16, 8, 7), (16, 8, 6), (16, 8, 5), (16, 4, 4), (16, 6, 8), (12, 8, 5),
              (12, 8, 6), (12, 8, 7), (12, 8, 8), (12, 8, 9), (12, 8, 10), (12, 8, 12), (10, 8, 16),
              (12, 12, 4), (12, 10, 10), (8, 8, 8)]
    print("QFLEX_TC_DEBIAS=%s" % os.environ.get("QFLEX_TC_DEBIAS", "(default: on)"), flush=True)
    for dist in ("uniform", "normal"):
        for (m, n, k) in shapes:
            a_k = list(range(k)); a_m = list(range(k, k + m))
            b_n = list(range(n)); b_k = list(range(n, n + k))
            c_n = list(range(n)); c_m = list(range(n, n + m))
            sa, sb, sc = 1 << (m + k), 1 << (n + k), 1 << (m + n)
            if dist == "uniform":
                a = (rng.uniform(-1, 1, sa) + 1j * rng.uniform(-1, 1, sa)).astype(np.complex64)
                b = (rng.uniform(-1, 1, sb) + 1j * rng.uniform(-1, 1, sb)).astype(np.complex64)
            else:
                a = rng.standard_normal(2 * sa, dtype=np.float32).view(np.complex64)
                b = rng.standard_normal(2 * sb, dtype=np.float32).view(np.complex64)
            ctx.reserve(sa + sb + sc + 128)
            ctx.upload(0, a); ctx.upload(sa + 32, b)
            rows = np.sort(rng.choice(1 << m, size=min(1 << m, 512), replace=False))
            A = a.reshape(1 << m, 1 << k)[rows].astype(np.complex128)
            want = A @ b.reshape(1 << k, 1 << n).astype(np.complex128)
            row = dict(dist=dist, m=m, n=n, k=k)
            for name, prec in (("tc", 0), ("ffma", 1)):
                before = ctx.stats()
                ctx.contract(_device.make_desc(0, sa + 32, sa + sb + 64, a_m, c_m, b_n, c_n, a_k, b_k,
                                               precision=prec))
                after = ctx.stats()
                got = ctx.download(sa + sb + 64, sc).reshape(1 << m, 1 << n)[rows].astype(np.complex128)
                d = got - want
                row[name + "_err"] = float(np.linalg.norm(d) / np.linalg.norm(want))
                row[name + "_bias"] = float(np.real(np.vdot(want, d)) / np.real(np.vdot(want, want)))
                if name == "tc":
                    row["kernel"] = ("pair" if after["launches_gemm_tc_pair"] > before["launches_gemm_tc_pair"]
                                     else "fused" if after["launches_gemm_tc_fused"] > before["launches_gemm_tc_fused"]
                                     else "tc" if after["launches_gemm_tc"] > before["launches_gemm_tc"] else "ffma")
            print(json.dumps(row), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
