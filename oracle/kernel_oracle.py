"""TEST INFRASTRUCTURE ONLY (oracle/).  numpy restatement of the three tensor
primitives on qFlex's hot path, in the bit-addressed form of
include/qflex_b200.h.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline leg may import this; nothing under qflex_b200/ does.

  contract  : qflex::multiply (reference src/tensor.cpp:1045-1194) =
              reorder A -> [left, common], reorder B -> [common, right]
              (src/tensor.cpp:1102-1121), cblas_cgemm RowMajor N,N alpha=1 beta=0
              (src/tensor.cpp:960-980; cgemv/cdotu are its N=1 / M=N=1 cases,
              :982-1042), C indices = left ++ right (:1167-1182).
  permute   : Tensor::reorder (src/tensor.cpp:866-911) restricted to
              power-of-two dims = bit permutation of the linear offset, the same
              statement the reference's own test oracle `simple_reordering` uses
              (tests/src/tensor_test.cpp:44-74).
  multiply_named : multiply() on named-index tensors exactly as the reference
              defines left/right/common (src/tensor.cpp:1065-1087).

Pinned by tests/test_oracle.py against the reference's golden vectors
(tensor_test.cpp:496-530 Multiply = {20,14,100,78}; reorder tests :234-469 via
the compiled reference in oracle/_ref) -- see DESIGN.md "Oracle".
"""
import numpy as np


def deposit(x, pos):
    """Bit i of x -> bit pos[i] of the result (vectorised over x)."""
    x = np.asarray(x, dtype=np.uint64)
    out = np.zeros_like(x)
    for i, p in enumerate(pos):
        out |= ((x >> np.uint64(i)) & np.uint64(1)) << np.uint64(p)
    return out.astype(np.int64)


def contract(a, b, a_m, c_m, b_n, c_n, a_k, b_k, c_size=None, acc_dtype=np.complex64):
    """C[dep(m,c_m)+dep(n,c_n)] = sum_k A[dep(m,a_m)+dep(k,a_k)] * B[dep(k,b_k)+dep(n,b_n)].

    a, b: flat complex64 arrays (the operand spans).  Returns flat C.
    """
    M, N, K = 1 << len(a_m), 1 << len(b_n), 1 << len(a_k)
    ma, mc = deposit(np.arange(M), a_m), deposit(np.arange(M), c_m)
    nb, nc = deposit(np.arange(N), b_n), deposit(np.arange(N), c_n)
    ka, kb = deposit(np.arange(K), a_k), deposit(np.arange(K), b_k)
    a2 = np.asarray(a)[ma[:, None] + ka[None, :]].astype(acc_dtype)
    b2 = np.asarray(b)[kb[:, None] + nb[None, :]].astype(acc_dtype)
    c2 = (a2 @ b2).astype(np.complex64)
    if c_size is None:
        c_size = M * N
    c = np.zeros(c_size, dtype=np.complex64)
    c[mc[:, None] + nc[None, :]] = c2
    return c


def dot(a, b, a_k, b_k):
    """Scalar sum_k A[dep(k,a_k)] * B[dep(k,b_k)] in complex128 (reference:
    cdotu then `amplitude += complex<double>(result)`, contraction_utils.cpp:352)."""
    K = 1 << len(a_k)
    ka, kb = deposit(np.arange(K), a_k), deposit(np.arange(K), b_k)
    return np.sum(np.asarray(a)[ka].astype(np.complex128) * np.asarray(b)[kb].astype(np.complex128))


def permute(src, dst_pos):
    """dst[sigma(i)] = src[i], bit b of i -> bit dst_pos[b] of sigma(i)."""
    n = 1 << len(dst_pos)
    out = np.empty(n, dtype=np.asarray(src).dtype)
    out[deposit(np.arange(n), dst_pos)] = np.asarray(src)[:n]
    return out


def reorder_named(data, indices, dims, new_indices):
    """Tensor::reorder on a named-index tensor (any dims): plain transpose."""
    t = np.asarray(data).reshape(dims)
    perm = [indices.index(x) for x in new_indices]
    return np.ascontiguousarray(t.transpose(perm)).reshape(-1), [dims[p] for p in perm]


def multiply_named(a, a_idx, a_dims, b, b_idx, b_dims):
    """qflex::multiply: returns (C flat, C indices, C dims)."""
    left = [x for x in a_idx if x not in b_idx]
    right = [x for x in b_idx if x not in a_idx]
    common = [x for x in a_idx if x in b_idx]
    a2, _ = reorder_named(a, a_idx, a_dims, left + common)
    b2, _ = reorder_named(b, b_idx, b_dims, common + right)
    dl = int(np.prod([a_dims[a_idx.index(x)] for x in left], dtype=np.int64))
    dr = int(np.prod([b_dims[b_idx.index(x)] for x in right], dtype=np.int64))
    dc = int(np.prod([a_dims[a_idx.index(x)] for x in common], dtype=np.int64))
    c = (a2.reshape(dl, dc).astype(np.complex64) @ b2.reshape(dc, dr).astype(np.complex64))
    return (c.reshape(-1).astype(np.complex64), left + right,
            [a_dims[a_idx.index(x)] for x in left] + [b_dims[b_idx.index(x)] for x in right])


def bits_from_named(a_idx, a_dims, b_idx, b_dims):
    """Named power-of-two tensors -> the bit tables of qfx_contract_desc
    (same construction as the host planner, qflex_b200/csrc/host/plan.cpp)."""
    def layout(idx, dims):
        pos, p = {}, 0
        for name, d in reversed(list(zip(idx, dims))):
            w = int(d).bit_length() - 1
            assert 1 << w == d, "power-of-two dims only"
            pos[name] = (p, w)
            p += w
        return pos
    left = [x for x in a_idx if x not in b_idx]
    right = [x for x in b_idx if x not in a_idx]
    common = [x for x in a_idx if x in b_idx]
    pa, pb = layout(a_idx, a_dims), layout(b_idx, b_dims)
    c_idx = left + right
    c_dims = [a_dims[a_idx.index(x)] for x in left] + [b_dims[b_idx.index(x)] for x in right]
    pc = layout(c_idx, c_dims)
    t = dict(a_m=[], c_m=[], b_n=[], c_n=[], a_k=[], b_k=[])
    for x in reversed(left):
        for bit in range(pa[x][1]):
            t["a_m"].append(pa[x][0] + bit)
            t["c_m"].append(pc[x][0] + bit)
    for x in reversed(right):
        for bit in range(pb[x][1]):
            t["b_n"].append(pb[x][0] + bit)
            t["c_n"].append(pc[x][0] + bit)
    for x in reversed(common):
        for bit in range(pa[x][1]):
            t["a_k"].append(pa[x][0] + bit)
            t["b_k"].append(pb[x][0] + bit)
    return t, c_idx, c_dims
