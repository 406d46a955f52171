"""Normwise error of the tcgen05 split-bf16 path vs FFMA vs complex128 oracle."""
import os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from qflex_b200 import _device
from oracle import kernel_oracle as ko

def main():
    ctx = _device.Context(0)
    rng = np.random.default_rng(0)
    for (m, n, k) in [(10, 7, 5), (10, 7, 6), (10, 7, 8), (10, 7, 10), (9, 7, 12), (8, 6, 14), (7, 5, 16)]:
        a_k = list(range(k)); a_m = list(range(k, k + m))
        b_n = list(range(n)); b_k = list(range(n, n + k))
        c_n = list(range(n)); c_m = list(range(n, n + m))
        sa, sb, sc = 1 << (m + k), 1 << (n + k), 1 << (m + n)
        a = (rng.uniform(-1, 1, sa) + 1j * rng.uniform(-1, 1, sa)).astype(np.complex64)
        b = (rng.uniform(-1, 1, sb) + 1j * rng.uniform(-1, 1, sb)).astype(np.complex64)
        ctx.reserve(sa + sb + sc + 128)
        ctx.upload(0, a); ctx.upload(sa + 32, b)
        want = ko.contract(a, b, a_m, c_m, b_n, c_n, a_k, b_k, acc_dtype=np.complex128)
        row = dict(m=m, n=n, k=k)
        for name, prec in (("tc", 0), ("ffma", 1)):
            ctx.contract(_device.make_desc(0, sa + 32, sa + sb + 64, a_m, c_m, b_n, c_n, a_k, b_k, precision=prec))
            got = ctx.download(sa + sb + 64, sc)
            row[name] = float(np.linalg.norm(got - want) / np.linalg.norm(want))
        row["f32"] = float(np.linalg.norm(ko.contract(a, b, a_m, c_m, b_n, c_n, a_k, b_k) - want) / np.linalg.norm(want))
        print(json.dumps(row), flush=True)
    ctx.close()

if __name__ == "__main__":
    main()
