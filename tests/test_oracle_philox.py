"""Pins the oracle's Philox4x32-10 to the Random123 known-answer vectors and checks the
sampling layout the reference defines (BaseModel.py:332-408, ns_strategy.py, utils.py:11-16)."""
import numpy as np
import pytest
from oracle import philox

KAT = [  # Random123 kat_vectors: philox4x32 10  ctr[4] key[2] -> out[4]
    ([0, 0, 0, 0], [0, 0], [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
    ([0xffffffff] * 4, [0xffffffff] * 2, [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
    ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0],
     [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]),
]


@pytest.mark.parametrize("ctr,key,out", KAT)
def test_philox_known_answers(ctr, key, out):
    assert philox.philox4x32_10(ctr, key).tolist() == out


def test_uniform_shape_range_and_determinism():
    X = np.arange(30).reshape(10, 3)
    a, k = philox.negatives(5, 0, X, 4, "t", 100)
    b, _ = philox.negatives(5, 0, X, 4, "t", 100)
    c, _ = philox.negatives(5, 1, X, 4, "t", 100)
    assert a.shape == (40,) and k == 4 and a.min() >= 0 and a.max() < 100
    assert np.array_equal(a, b) and not np.array_equal(a, c)


def test_ht_interleaves_sides_and_halves_ratio():
    X = np.arange(30).reshape(10, 3)
    ids, k = philox.negatives(1, 0, X, 5, "h+t", 50)        # 5 // 2 = 2 per side
    assert k == 4 and ids.shape == (40,)
    h = philox.sample_uniform(1, 0, "h", 10, 2, 50).reshape(10, 2)
    t = philox.sample_uniform(1, 0, "t", 10, 2, 50).reshape(10, 2)
    got = ids.reshape(10, 2, 2)
    assert np.array_equal(got[:, :, 0], h) and np.array_equal(got[:, :, 1], t)


def test_uniform_pool_lookup():
    pool = np.array([7, 9, 11])
    ids = philox.sample_uniform(3, 0, "h", 50, 3, 1000, pool)
    assert set(ids.tolist()) <= {7, 9, 11}


def test_typed_keeps_type_and_excludes_self():
    """tests/test_ns_strategy.py:52-53,61-62: sampled entity has the replaced entity's type."""
    rng = np.random.default_rng(0)
    ind2type = rng.choice(["A", "B", "C"], 60).tolist()
    pools = philox.TypedPools(ind2type)
    X = np.stack([rng.integers(0, 60, 200), rng.integers(0, 5, 200), rng.integers(0, 60, 200)], 1)
    for side, col in (("h", 0), ("t", 2)):
        ids = philox.sample_typed(9, 2, side, X, 3, pools)
        ref = np.repeat(X[:, col], 3)
        assert all(ind2type[a] == ind2type[b] for a, b in zip(ids, ref))
        assert (ids != ref).all()


def test_typed_matches_np_delete_form():
    ind2type = ["A", "B", "A", "A", "B", "A"]
    pools = philox.TypedPools(ind2type)
    X = np.array([[0, 0, 1], [3, 0, 4], [5, 0, 2]])
    u = philox.draw_u32(4, 7, 0, 6).astype(np.int64)
    ids = philox.sample_typed(4, 7, "h", X, 2, pools)
    for n, (x, got) in enumerate(zip(np.repeat(X[:, 0], 2), ids)):
        pool = np.array([i for i, t in enumerate(ind2type) if t == ind2type[x]])
        pool = np.delete(pool, np.where(pool == x))           # utils.py:12-13
        assert got == pool[u[n] % len(pool)]


def test_typed_singleton_pool_raises():
    pools = philox.TypedPools(["A", "B", "B"])
    with pytest.raises(ValueError):
        philox.sample_typed(1, 0, "h", np.array([[0, 0, 1]]), 2, pools)


def test_epoch_permutation_is_a_reproducible_permutation():
    """Spec of kge_epoch_perm (csrc/batch_stream.cu): stable argsort of the Philox keys of (seed, pass)."""
    import numpy as np
    from oracle import philox
    a = philox.epoch_permutation(7, 0, 5000)
    assert np.array_equal(np.sort(a), np.arange(5000)) and a.dtype == np.uint32
    assert np.array_equal(a, philox.epoch_permutation(7, 0, 5000))
    assert not np.array_equal(a, philox.epoch_permutation(7, 1, 5000)) and not np.array_equal(a, philox.epoch_permutation(8, 0, 5000))
