"""Device-side known-positive filter index (kge_filter_index_build / kge_filter_lookup)
against the reference's rule — positive_X[(positive_X[:,1] == r) & (positive_X[:,fs] == x[fs]), cs]
(BaseModel.py:646-650) — restated as numpy masks, and through kge_rank in both filter forms."""
import numpy as np
import pytest
import torch

from helpers import make_pair, rand_triples, filtered_ids
from oracle import ref_ops as O

pytestmark = pytest.mark.gpu


def _sets(be, ents):
    be, ents = be.cpu().numpy(), ents.cpu().numpy()
    return [ents[b:e] for b, e in zip(be[0], be[1])]


@pytest.mark.parametrize("side", ["h", "t"])
@pytest.mark.parametrize("N,R,n_pos", [(40, 5, 300), (257, 9, 4000), (14951, 1345, 60000)])
def test_index_matches_reference_masks(side, N, R, n_pos):
    from KGE._b200.filter import build_filter
    rng = np.random.default_rng(N + n_pos)
    pos = rand_triples(rng, n_pos, N, R)
    pos[: n_pos // 10] = pos[n_pos // 10: 2 * (n_pos // 10)]              # duplicated triples: listed once
    Q = pos[rng.integers(0, n_pos, 200)].copy()
    Q[::3] = rand_triples(rng, len(Q[::3]), N, R)                          # queries with empty filters
    be, ents = build_filter(pos, Q, side, N, "cuda:0", n_rel=R)
    got = _sets(be, ents)
    for q, x in enumerate(Q):
        want = np.unique(filtered_ids(pos, x, side))
        assert np.array_equal(got[q], want), (side, q, got[q], want)       # sorted, unique, complete


def test_index_edge_cases():
    from KGE._b200.filter import build_filter, FilterIndex
    from KGE._b200 import engine as E, lib as L
    N, R = 50, 4
    # ids outside the tables are dropped, not matched; a single triple; the largest ids
    pos = np.array([[49, 3, 49], [0, 0, 0], [49, 3, 48], [50, 0, 1], [1, 4, 1], [-1, 0, 2], [49, 3, 49]], dtype=np.int64)
    Q = np.array([[49, 3, 7], [0, 0, 0], [7, 3, 49], [1, 0, 1]], dtype=np.int64)
    be, ents = build_filter(pos, Q, "t", N, "cuda:0", n_rel=R)
    assert [s.tolist() for s in _sets(be, ents)] == [[48, 49], [0], [], []]
    be, ents = build_filter(pos, Q, "h", N, "cuda:0", n_rel=R)
    assert [s.tolist() for s in _sets(be, ents)] == [[], [0], [49], []]
    cfg = L.Cfg(abi_version=L.KGE_ABI_VERSION, model=L.TRANSE, d_ent=1, d_rel=1, id_dtype=1, score_p=1.0, n_ent=N, n_rel=R)
    idx = FilterIndex(E._ctx(0), cfg, pos, "t", "cuda:0")
    assert idx.n_unique() == 3                                             # [49,3,49] x2, [0,0,0], [49,3,48]
    with pytest.raises(ValueError):
        FilterIndex(E._ctx(0), cfg, np.zeros((0, 3), np.int64), "t", "cuda:0")
    # int32 ids (what the CSV iterators yield, tests/test_data_utils.py:82,90)
    cfg32 = L.Cfg(abi_version=L.KGE_ABI_VERSION, model=L.TRANSE, d_ent=1, d_rel=1, id_dtype=0, score_p=1.0, n_ent=N, n_rel=R)
    idx32 = FilterIndex(E._ctx(0), cfg32, pos.astype(np.int32), "t", "cuda:0")
    be32, ents32 = idx32.lookup(torch.as_tensor(Q.astype(np.int32)).cuda())
    assert [s.tolist() for s in _sets(be32, ents32)] == [[48, 49], [0], [], []]


@pytest.mark.parametrize("model,score,d", [("TransE", O.LpDistance(1), 50), ("DistMult", None, 64), ("TransH", O.LpDistancePow(2), 24),
                                           ("TransR", O.LpDistancePow(2), 16)])
def test_range_filter_equals_packed_csr_through_kge_rank(model, score, d):
    """kge_filter.end (ranges into the index) and the packed CSR form give the same counts on every
    ranking path (tile, tensor core, generic, matrix)."""
    from KGE._b200.filter import build_filter
    N, R = 700, 9
    rng = np.random.default_rng(8)
    _, _, eng = make_pair(model, N, R, d, score=score, loss=O.PairwiseHingeLoss(1.0))
    pos = rand_triples(rng, 30000, N, R)
    Q = pos[rng.integers(0, len(pos), 300)].copy()
    for side, s in (("h", 0), ("t", 1)):
        be, ents = build_filter(pos, Q, side, N, eng.device, n_rel=R)
        lens = (be[1] - be[0])
        off = torch.zeros(len(Q) + 1, dtype=torch.int64, device=eng.device)
        off[1:] = torch.cumsum(lens, 0)
        packed = torch.cat([ents[b:e] for b, e in zip(be[0].tolist(), be[1].tolist())] + [ents[:0]])
        a = eng.rank_counts(Q, s, be, ents)
        b = eng.rank_counts(Q, s, off, packed if packed.numel() else ents[:1])
        raw = eng.rank_counts(Q, s)
        assert torch.equal(a, b) and bool((a <= raw).all()) and bool((a < raw).any())
