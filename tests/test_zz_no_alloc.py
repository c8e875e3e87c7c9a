"""kge_rank_lp_static (include/kge_abi.h): the allocation-free variant of kge_rank's Lp-tile path, whose
scratch is a static __device__ array inside the library.  Through the C-ABI, against kge_rank on the same
inputs, two calls in a row (the array keeps the previous call's contents), plus the size limit.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest
import torch

from helpers import make_pair, rand_triples
from oracle import ref_ops as O

pytestmark = pytest.mark.gpu


def _engine(model, score, d, N=1500, R=9):
    _, _, eng = make_pair(model, N, R, d, score=score, loss=O.PairwiseHingeLoss(1.0))
    return eng


@pytest.mark.parametrize("model,score,d", [("TransE", O.LpDistance(1), 50), ("RotatE", O.LpDistance(1), 21),
                                           ("UM", O.LpDistancePow(2), 33)])
def test_static_scratch_rank_equals_kge_rank_twice(model, score, d):
    from KGE._b200.filter import build_filter
    N, R = 1500, 9
    eng = _engine(model, score, d, N, R)
    rng = np.random.default_rng(31)
    pos = rand_triples(rng, 20000, N, R)
    # two DIFFERENT calls in a row: the second, smaller one runs over the first one's left-overs
    for Q, side in ((700, "t"), (130, "h")):
        Xq = pos[rng.integers(0, len(pos), Q)].copy()
        Xq[::4] = rand_triples(rng, len(Xq[::4]), N, R)
        off, ids = build_filter(pos, Xq, side, N, eng.device)
        s = 0 if side == "h" else 1
        for flt in ((None, None), (off, ids)):
            want = eng.rank_counts(Xq, s, *flt)
            got = eng.rank_counts_static(Xq, s, *flt)
            # same kernels, same arithmetic, integer counts: bit-exact (the tests' bar for integer work)
            assert torch.equal(got, want), (model, Q, side, int((got != want).sum()))
        # entity shards add up through the static entry point too
        parts = sum(eng.rank_counts_static(Xq, s, off, ids, ent_lo=lo, ent_hi=hi) for lo, hi in ((0, 500), (500, 501), (501, N)))
        assert torch.equal(parts, eng.rank_counts(Xq, s, off, ids))


def test_static_scratch_rejects_what_does_not_fit_and_other_paths():
    from KGE._b200 import lib as L
    eng = _engine("TransE", O.LpDistance(1), 50)
    per_q = 4 * (52 + 1)
    too_many = L.RANK_STATIC_BYTES // per_q + 1024
    Xq = rand_triples(np.random.default_rng(1), too_many, 1500, 9)
    with pytest.raises(RuntimeError, match="static array holds"):
        eng.rank_counts_static(Xq, 1)
    ok = rand_triples(np.random.default_rng(2), 19712, 1500, 9)          # the documented C1 maximum fits
    assert torch.equal(eng.rank_counts_static(ok, 1), eng.rank_counts(ok, 1))
    _, _, dm = make_pair("DistMult", 300, 5, 16, loss=O.BinaryCrossEntropyLoss())
    with pytest.raises(RuntimeError, match="Lp-tile path only"):
        dm.rank_counts_static(rand_triples(np.random.default_rng(3), 8, 300, 5), 1)


def test_static_entry_point_allocates_nothing():
    """torch's caching allocator sees no request and the driver's free memory does not move across calls
    whose output buffer is pre-allocated (ids already on the device, output passed in)."""
    eng = _engine("TransE", O.LpDistance(1), 50)
    Xq = torch.as_tensor(rand_triples(np.random.default_rng(4), 512, 1500, 9)).to(eng.device)
    out = torch.zeros(512, dtype=torch.int32, device=eng.device)
    eng.rank_counts_static(Xq, 1, out=out)                                # first call resolves the symbol address
    torch.cuda.synchronize()
    n0 = torch.cuda.memory_stats()["allocation.all.allocated"]
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(3):
        out.zero_()
        eng.rank_counts_static(Xq, 1, out=out)
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats()["allocation.all.allocated"] == n0
    assert torch.cuda.mem_get_info()[0] == free0
    assert torch.equal(out, eng.rank_counts(Xq, 1))
