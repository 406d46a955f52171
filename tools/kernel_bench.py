"""Kernel micro-benchmarks through the C ABI (CUDA events on the context stream).
Shapes = the distinct (M,N,K) classes of SURVEY Appendix A."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from qflex_b200 import _device  # noqa: E402


def layout(m, n, k, mode):
    """mode 'gemm': A=[m|k], B=[k|n]; 'mixed': K bits interleaved inside A."""
    if mode == "gemm":
        a_k = list(range(k)); a_m = list(range(k, k + m))
    else:  # K split in two groups: low group at bits [4, 4+k/2), high group in the middle
        h = k // 2
        pos = list(range(m + k))
        lo = pos[min(4, m):min(4, m) + h]
        mid = m // 2 + h
        hi = pos[min(4, m) + h + mid - h: min(4, m) + h + mid - h + (k - h)]
        a_k = lo + hi
        a_m = [p for p in pos if p not in a_k]
    if os.environ.get("QFLEX_KB_MLOW"):  # output bits lowest in A (open-output tails)
        a_m = list(range(m)); a_k = list(range(m, m + k))
    elif os.environ.get("QFLEX_KB_MMID"):  # output bits in the middle of A's address
        a_m = list(range(6, 6 + m)); a_k = list(range(6)) + list(range(6 + m, m + k))
    b_n = list(range(n)); b_k = list(range(n, n + k))
    c_n = list(range(n)); c_m = list(range(n, n + m))
    if os.environ.get("QFLEX_KB_CT"):  # C transposed: m lowest (what a role swap produces)
        c_m = list(range(m)); c_n = list(range(m, m + n))
    return a_m, c_m, b_n, c_n, a_k, b_k


def main():
    shapes = [(20, 8, 8), (16, 8, 8), (15, 6, 6), (12, 6, 6), (24, 4, 4), (20, 5, 5), (18, 3, 3),
              (12, 8, 16), (9, 6, 12), (12, 12, 12), (22, 1, 1), (22, 2, 1)]
    if len(sys.argv) > 1:
        shapes = [tuple(int(x) for x in s.split(",")) for s in sys.argv[1:]]
    ctx = _device.Context(0)
    rng = np.random.default_rng(0)
    out = []
    for (m, n, k) in shapes:
        for mode in (("gemm",) if (n == 0 or m == 0) else ("gemm", "mixed")):
            a_m, c_m, b_n, c_n, a_k, b_k = layout(m, n, k, mode)
            sa, sb, sc = 1 << (m + k), 1 << (n + k), 1 << (m + n)
            ctx.reserve(sa + sb + sc + 128)
            blk = (rng.uniform(-1, 1, 1 << 16) + 1j * rng.uniform(-1, 1, 1 << 16)).astype(np.complex64)
            for off, size in ((0, sa), (sa + 32, sb)):
                for o in range(0, size, 1 << 16):
                    ctx.upload(off + o, blk[: min(1 << 16, size - o)])
                if size > (1 << 22):
                    break  # big A: first 2^16 random, rest whatever the arena holds (finite)
            d = _device.make_desc(0, sa + 32, sa + sb + 64, a_m, c_m, b_n, c_n, a_k, b_k,
                                  precision=int(os.environ.get("QFLEX_KB_PREC", "0")))
            for _ in range(2):
                ctx.contract(d)
            ctx.sync()
            reps = 20 if m + n + k < 26 else 5
            ctx.timer_start()
            for _ in range(reps):
                ctx.contract(d)
            ms = ctx.timer_stop_ms() / reps
            flops = 8.0 * 2 ** (m + n + k)
            byts = 8.0 * (sa + sb + sc)
            row = dict(m=m, n=n, k=k, mode=mode, ms=round(ms, 4), tflops=round(flops / ms / 1e9, 2),
                       gbs=round(byts / ms / 1e6, 1))
            print(json.dumps(row), flush=True)
            out.append(row)
    # permute
    for nbits in (() if os.environ.get("QFLEX_KB_NOPERM") else (24, 28)):
        n = 1 << nbits
        ctx.reserve(2 * n + 64)
        for name, p in (("reverse", list(reversed(range(nbits)))),
                        ("rot8", [(b + 8) % nbits for b in range(nbits)]),
                        ("identity", list(range(nbits)))):
            for _ in range(2):
                ctx.permute(0, n + 32, p)
            ctx.sync()
            ctx.timer_start()
            for _ in range(5):
                ctx.permute(0, n + 32, p)
            ms = ctx.timer_stop_ms() / 5
            row = dict(permute=name, nbits=nbits, ms=round(ms, 4), gbs=round(16.0 * n / ms / 1e6, 1))
            print(json.dumps(row), flush=True)
            out.append(row)
    ctx.close()


if __name__ == "__main__":
    main()
