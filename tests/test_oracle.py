"""The oracle (oracle/kernel_oracle.py) pinned against the reference's golden
vectors and against the compiled reference itself (oracle/_ref)."""
import numpy as np
import pytest

from oracle import kernel_oracle as ko
from tests import support


def _rand(n, rng):
    return (rng.uniform(-1, 1, n) + 1j * rng.uniform(-1, 1, n)).astype(np.complex64)


def test_multiply_golden_tensor_test():
    # reference tests/src/tensor_test.cpp:496-530
    a = np.arange(8).astype(np.complex64)
    b = np.arange(8, 0, -1).astype(np.complex64)
    c, idx, dims = ko.multiply_named(a, ["a", "b", "c"], [2, 2, 2], b, ["b", "c", "d"], [2, 2, 2])
    assert idx == ["a", "d"] and dims == [2, 2]
    np.testing.assert_allclose(c.real, [20, 14, 100, 78])
    t, cidx, cdims = ko.bits_from_named(["a", "b", "c"], [2, 2, 2], ["b", "c", "d"], [2, 2, 2])
    np.testing.assert_allclose(ko.contract(a, b, **t).real, [20, 14, 100, 78])
    assert cidx == ["a", "d"] and cdims == [2, 2]


def test_bit_form_equals_named_form():
    rng = np.random.default_rng(1234)
    a_idx, a_dims = ["p", "q", "r", "s", "t"], [4, 2, 8, 2, 4]
    b_idx, b_dims = ["t", "u", "q", "r", "v"], [4, 4, 2, 8, 2]
    a, b = _rand(int(np.prod(a_dims)), rng), _rand(int(np.prod(b_dims)), rng)
    want, idx, dims = ko.multiply_named(a, a_idx, a_dims, b, b_idx, b_dims)
    t, cidx, cdims = ko.bits_from_named(a_idx, a_dims, b_idx, b_dims)
    got = ko.contract(a, b, **t)
    assert cidx == idx and cdims == dims
    assert support.rel_err(got, want) < 1e-6


@pytest.mark.skipif(not support.reference_available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("seed", [0, 1, 2])
def test_oracle_multiply_matches_reference(seed):
    rng = np.random.default_rng(seed)
    letters = list("abcdefgh")
    rng.shuffle(letters)
    a_idx = letters[:5]
    b_idx = [letters[i] for i in rng.permutation(5)[:3]] + letters[5:7]
    rng.shuffle(b_idx)
    dim_of = {l: int(2 ** rng.integers(1, 4)) for l in letters}
    a_dims, b_dims = [dim_of[x] for x in a_idx], [dim_of[x] for x in b_idx]
    a, b = _rand(int(np.prod(a_dims)), rng), _rand(int(np.prod(b_dims)), rng)
    want, widx, wdims = support.reference_multiply(a, a_idx, a_dims, b, b_idx, b_dims)
    got, idx, dims = ko.multiply_named(a, a_idx, a_dims, b, b_idx, b_dims)
    assert idx == widx and dims == wdims
    assert support.rel_err(got, want) < 1e-5
    t, _, _ = ko.bits_from_named(a_idx, a_dims, b_idx, b_dims)
    assert support.rel_err(ko.contract(a, b, **t), want) < 1e-5


@pytest.mark.skipif(not support.reference_available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("nbits,seed", [(3, 0), (10, 1), (12, 2), (12, 3), (14, 4), (16, 5)])
def test_oracle_permute_matches_reference_reorder(nbits, seed):
    # the reorder patterns of tensor_test.cpp:234-469 (right-only, left-only,
    # L-R-L worst case) are all bit permutations; random ones cover them
    rng = np.random.default_rng(seed)
    idx = ["i%d" % i for i in range(nbits)]
    dims = [2] * nbits
    perm = list(rng.permutation(nbits))
    new_idx = [idx[p] for p in perm]
    data = _rand(1 << nbits, rng)
    want = support.reference_reorder(data, idx, dims, new_idx)
    # index at position q (from the left) owns bit nbits-1-q
    dst_pos = [0] * nbits
    for new_q, old_q in enumerate(perm):
        dst_pos[nbits - 1 - old_q] = nbits - 1 - new_q
    got = ko.permute(data, dst_pos)
    assert np.array_equal(got, want)
    named, _ = ko.reorder_named(data, idx, dims, new_idx)
    assert np.array_equal(named, want)


GENERAL_REORDERS = [([3, 2, 5], [2, 0, 1]), ([2, 4, 3, 1], [3, 1, 0, 2]), ([6, 7], [1, 0]),
                    ([5, 1, 2, 3, 2], [4, 2, 0, 3, 1]), ([3, 3, 3, 3], [0, 1, 2, 3]),
                    ([4, 3, 256], [1, 2, 0]), ([3, 1024], [1, 0])]


@pytest.mark.skipif(not support.reference_available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("dims,order", GENERAL_REORDERS)
def test_oracle_reorder_general_dimensions_matches_reference(dims, order):
    """Non-power-of-two dimensions take Tensor::_naive_reorder in the reference
    (src/tensor.cpp:377-492); the oracle's transpose is pinned to it here, and the CUDA
    qfx_reorder_general to the oracle in tests/test_gpu_kernels.py.  Sizes are <= 1024
    or multiples of 1024: the reference walks the tensor in blocks of MAX_RIGHT_DIM = 1024
    and copies a full block even when fewer elements remain (src/tensor.cpp:474-478), so
    e.g. {12, 10, 9} reads past its scratch and is not a usable golden."""
    rng = np.random.default_rng(sum(dims))
    idx = ["a%d" % i for i in range(len(dims))]
    new_idx = [idx[p] for p in order]
    data = _rand(int(np.prod(dims)), rng)
    want = support.reference_reorder(data, idx, dims, new_idx)
    got, new_dims = ko.reorder_named(data, idx, dims, new_idx)
    assert new_dims == [dims[p] for p in order]
    assert np.array_equal(got, want)
