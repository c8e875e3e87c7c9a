"""GPU parity of filtered ranking (kge_rank) against the oracle's get_rank
(BaseModel.py:620-654).  Ranks must match exactly, except where the oracle's own
candidate score is within 1e-5 relative of the true triple's score (documented ties)."""
import numpy as np
import pytest
import torch

from helpers import make_pair, rand_triples, RankBand, oracle_rank_check
from oracle import ref_ops as O, ref_step as RS
from test_gpu_step import SCORES, LOSSES, VECTOR_MODELS, _score_for

pytestmark = pytest.mark.gpu
N, R = 257, 9


def _check(model, score, side, filtered, d=50, d_rel=None, Q=40):
    from KGE._b200.filter import build_filter
    rng = np.random.default_rng(2)
    ref, W, eng = make_pair(model, N, R, d, d_rel, score=_score_for(model, score), loss=LOSSES["hinge"])
    pos = rand_triples(rng, 4000, N, R)
    Xq = pos[:Q].copy()
    Xq[::3] = rand_triples(rng, len(Xq[::3]), N, R)        # some queries not in positive_X
    off = ids = None
    if filtered:
        off, ids = build_filter(pos, Xq, side, N, eng.device)
    cnt = eng.rank_counts(Xq, 0 if side == "h" else 1, off, ids).cpu().numpy()
    band = RankBand()
    for q, x in enumerate(Xq):
        oracle_rank_check(band, ref, W, x, side, pos if filtered else None, int(cnt[q]) + 1, (model, q))
    print(model, score, side, band.summary())
    assert band.used <= max(2, Q // 10), band.summary()     # the band is an exception, not the rule


@pytest.mark.parametrize("model", VECTOR_MODELS)
@pytest.mark.parametrize("side", ["h", "t"])
@pytest.mark.parametrize("filtered", [False, True])
def test_rank_matches_oracle(model, side, filtered):
    score = {"UM": "l2pow", "TransE": "l1", "TransH": "l2pow", "TransD": "l2pow", "RotatE": "l1",
             "DistMult": "dot"}[model]
    _check(model, score, side, filtered, d=50 if model != "RotatE" else 21)


@pytest.mark.parametrize("score", ["l2", "linf", "dot"])
def test_rank_transe_scores(score):
    _check("TransE", score, "t", True, d=33)


def test_filtered_rank_not_worse_than_raw():
    """tests/test_model.py:46-49 invariant."""
    from KGE._b200.filter import build_filter
    rng = np.random.default_rng(4)
    ref, W, eng = make_pair("TransE", N, R, 16, score=SCORES["l2"], loss=LOSSES["hinge"])
    pos = rand_triples(rng, 3000, N, R)
    off, ids = build_filter(pos, pos[:64], "h", N, eng.device)
    raw = eng.rank_counts(pos[:64], 0).cpu().numpy()
    flt = eng.rank_counts(pos[:64], 0, off, ids).cpu().numpy()
    assert (flt <= raw).all() and (flt >= 0).all()


def test_entity_shards_sum_to_full():
    rng = np.random.default_rng(6)
    ref, W, eng = make_pair("DistMult", N, R, 40, loss=LOSSES["bce"])
    Xq = rand_triples(rng, 33, N, R)
    full = eng.rank_counts(Xq, 1).cpu()
    parts = sum(eng.rank_counts(Xq, 1, ent_lo=lo, ent_hi=hi).cpu() for lo, hi in [(0, 100), (100, 101), (101, N)])
    assert torch.equal(full, parts)


@pytest.mark.parametrize("model,score,d,dr", [("SE", "l1", 24, 24), ("TransR", "l2pow", 40, 24), ("RESCAL", "dot", 32, 32),
                                              ("TransR", "l2pow", 160, 160)])
@pytest.mark.parametrize("side", ["h", "t"])
def test_rank_matrix_models(model, score, d, dr, side):
    _check(model, score if model != "RESCAL" else "dot", side, True, d=d, d_rel=dr, Q=16)


# ---- tensor-core (tcgen05) ranking path vs the generic CUDA-core path -------------------
@pytest.mark.parametrize("model,score,d,n_ent", [("DistMult", "dot", 200, 5000), ("DistMult", "dot", 50, 777),
                                                 ("RESCAL", "dot", 64, 3000), ("TransE", "dot", 100, 4097),
                                                 ("UM", "dot", 33, 1500), ("DistMult", "dot", 500, 2100)])
@pytest.mark.parametrize("side", ["h", "t"])
def test_tensor_core_rank_equals_generic(model, score, d, n_ent, side):
    import os
    from KGE._b200.filter import build_filter
    rng = np.random.default_rng(12)
    ref, W, eng = make_pair(model, n_ent, R, d, None, score=_score_for(model, score), loss=LOSSES["hinge"])
    pos = rand_triples(rng, 20000, n_ent, R)
    Xq = pos[:700].copy()
    Xq[::4] = rand_triples(rng, len(Xq[::4]), n_ent, R)
    off, ids = build_filter(pos, Xq, side, n_ent, eng.device)
    s = 0 if side == "h" else 1
    tc_raw = eng.rank_counts(Xq, s).cpu().numpy()
    tc_flt = eng.rank_counts(Xq, s, off, ids).cpu().numpy()
    os.environ["KGE_RANK_GENERIC"] = "1"
    try:
        g_raw = eng.rank_counts(Xq, s).cpu().numpy()
        g_flt = eng.rank_counts(Xq, s, off, ids).cpu().numpy()
    finally:
        os.environ.pop("KGE_RANK_GENERIC")
    for a, b in ((tc_raw, g_raw), (tc_flt, g_flt)):
        diff = np.abs(a.astype(np.int64) - b)
        # both are exact-fp32 rankings with different summation orders: only near-ties may move
        assert diff.max() <= 2 and (diff > 0).mean() <= 0.01, (diff.max(), (diff > 0).mean())
    assert (tc_flt <= tc_raw).all()
    # shards add up exactly
    parts = sum(eng.rank_counts(Xq, s, off, ids, ent_lo=lo, ent_hi=hi).cpu().numpy()
                for lo, hi in [(0, n_ent // 3), (n_ent // 3, n_ent // 3 + 77), (n_ent // 3 + 77, n_ent)])
    assert np.array_equal(parts, tc_flt)


@pytest.mark.parametrize("model,score,d", [("DistMult", "dot", 32), ("TransE", "dot", 20)])
def test_tensor_core_rank_one_call_over_several_query_blocks(model, score, d):
    """kge_rank takes any number of queries: beyond 16 384 the tensor-core path walks query blocks inside ONE call
    (table split once, the known-positive bitmask of block i+1 built on the side stream only after block i has
    finished with the buffer).  40 000 queries in one call == the same queries in calls of at most 16 384, exactly."""
    from KGE._b200.filter import build_filter
    n_ent, Q = 600, 40000
    rng = np.random.default_rng(31)
    _, _, eng = make_pair(model, n_ent, R, d, None, score=_score_for(model, score), loss=LOSSES["hinge"])
    pos = rand_triples(rng, 30000, n_ent, R)
    Xq = pos[rng.integers(0, len(pos), Q)].copy()
    Xq[::5] = rand_triples(rng, len(Xq[::5]), n_ent, R)
    for side, s in (("h", 0), ("t", 1)):
        off, ids = build_filter(pos, Xq, side, n_ent, eng.device)
        whole = eng.rank_counts(Xq, s, off, ids).cpu().numpy()
        whole_raw = eng.rank_counts(Xq, s).cpu().numpy()
        parts, parts_raw = [], []
        for lo in range(0, Q, 16384):
            o2, i2 = build_filter(pos, Xq[lo:lo + 16384], side, n_ent, eng.device)
            parts.append(eng.rank_counts(Xq[lo:lo + 16384], s, o2, i2).cpu().numpy())
            parts_raw.append(eng.rank_counts(Xq[lo:lo + 16384], s).cpu().numpy())
        assert np.array_equal(whole, np.concatenate(parts)), side
        assert np.array_equal(whole_raw, np.concatenate(parts_raw)), side
        assert (whole <= whole_raw).all() and (whole < whole_raw).any()


@pytest.mark.parametrize("n_proto", [2, 5])
def test_tensor_core_rank_with_mass_ties_falls_back_exactly(n_proto):
    """A table of only a few distinct rows: hundreds of candidates tie with the true score EXACTLY and every one of
    them is "uncertain" for the tensor-core pass.  Five distinct rows: a long list of pairs for the exact fix-up;
    two: the list (512 pairs per query) overflows and the block is re-ranked in fp32 as a whole.  Counts (strict >)
    must equal the generic path's bit for bit."""
    import os
    from KGE._b200.filter import build_filter
    n_ent, d = 3000, 64
    rng = np.random.default_rng(3)
    ref, W, eng = make_pair("DistMult", n_ent, R, d, None, score=None, loss=LOSSES["hinge"])
    proto = eng.tab[0].storage[:n_proto].clone()
    eng.tab[0].storage.copy_(proto[torch.from_numpy(rng.integers(0, n_proto, n_ent)).cuda()])
    eng.tab[0].touch()
    pos = rand_triples(rng, 5000, n_ent, R)
    Xq = pos[:300]
    for side, s in (("h", 0), ("t", 1)):
        off, ids = build_filter(pos, Xq, side, n_ent, eng.device)
        tc = eng.rank_counts(Xq, s, off, ids).cpu().numpy()
        os.environ["KGE_RANK_GENERIC"] = "1"
        try:
            gen = eng.rank_counts(Xq, s, off, ids).cpu().numpy()
        finally:
            os.environ.pop("KGE_RANK_GENERIC")
        assert np.array_equal(tc, gen), (side, np.abs(tc - gen).max())
        assert tc.max() > 100            # the ties are real: whole groups of candidates rank above


@pytest.mark.parametrize("model,score,d,n_ent,side", [
    ("TransE", "l1", 50, 1500, "t"), ("TransE", "l1", 50, 1500, "h"),
    ("TransE", "l2", 37, 700, "h"), ("TransE", "l2pow", 128, 333, "t"),
    ("TransE", "linf", 20, 500, "t"), ("TransE", "l3", 24, 400, "h"),
    ("UM", "l2pow", 64, 900, "t"), ("UM", "l1", 33, 900, "h"),
    ("RotatE", "l1", 37, 600, "t"), ("RotatE", "l1", 64, 600, "h"), ("RotatE", "l2", 20, 300, "h"),
    ("RotatE", "linf", 16, 300, "t"), ("RotatE", "l3", 12, 300, "t"),
])
def test_lp_tile_rank_equals_generic(model, score, d, n_ent, side):
    """CUDA-core tiled Lp ranking (csrc/rank_lp_tile.cuh) against the warp-per-candidate kernel:
    both are exact fp32 rankings with different evaluation orders, so only candidates within
    rounding of the true score may move; shards must add up exactly; filtered <= raw."""
    import os
    from KGE._b200.filter import build_filter
    rng = np.random.default_rng(21)
    ref, W, eng = make_pair(model, n_ent, R, d, None, score=_score_for(model, score), loss=LOSSES["hinge"])
    pos = rand_triples(rng, 20000, n_ent, R)
    Xq = pos[:700].copy()
    Xq[::4] = rand_triples(rng, len(Xq[::4]), n_ent, R)
    off, ids = build_filter(pos, Xq, side, n_ent, eng.device)
    s = 0 if side == "h" else 1
    t_raw = eng.rank_counts(Xq, s).cpu().numpy()
    t_flt = eng.rank_counts(Xq, s, off, ids).cpu().numpy()
    os.environ["KGE_RANK_GENERIC"] = "1"
    try:
        g_raw = eng.rank_counts(Xq, s).cpu().numpy()
        g_flt = eng.rank_counts(Xq, s, off, ids).cpu().numpy()
    finally:
        os.environ.pop("KGE_RANK_GENERIC")
    for a, b in ((t_raw, g_raw), (t_flt, g_flt)):
        diff = np.abs(a.astype(np.int64) - b)
        assert diff.max() <= 2 and (diff > 0).mean() <= 0.01, (diff.max(), (diff > 0).mean())
    assert (t_flt <= t_raw).all() and (t_flt >= 0).all()
    parts = sum(eng.rank_counts(Xq, s, off, ids, ent_lo=lo, ent_hi=hi).cpu().numpy()
                for lo, hi in [(0, n_ent // 3), (n_ent // 3, n_ent // 3 + 77), (n_ent // 3 + 77, n_ent)])
    assert np.array_equal(parts, t_flt)
    # the oracle's ranks on a sample (eligible candidates within 1e-5 relative of the true score excepted)
    band = RankBand()
    for qi in range(40):
        oracle_rank_check(band, ref, W, Xq[qi], side, pos, int(t_flt[qi]) + 1, qi)
    print(model, score, side, band.summary())
    assert band.used <= 4, band.summary()


@pytest.mark.parametrize("model,score,d,dr,n_ent,side", [
    ("TransH", "l2pow", 50, None, 1500, "t"), ("TransH", "l1", 37, None, 700, "h"),
    ("TransD", "l2pow", 50, 30, 1500, "h"), ("TransD", "l1", 64, 64, 900, "t"), ("TransD", "l2", 33, 41, 600, "t"),
    ("TransR", "l2pow", 40, 24, 1000, "t"), ("TransR", "l1", 24, 40, 700, "h"), ("TransR", "l2", 160, 160, 400, "h"),
    ("SE", "l1", 24, None, 900, "h"), ("SE", "l2", 50, None, 600, "t"), ("SE", "linf", 20, None, 500, "t"),
    ("TransH", "l3", 24, None, 400, "t"),
])
def test_projected_table_rank_equals_generic(model, score, d, dr, n_ent, side):
    """Per-relation projected candidate tables + the Lp tile kernel (csrc/rank_proj_tile.cuh) against the
    warp-per-(query, candidate) kernel: exact fp32 rankings in different evaluation orders, so only candidates
    within rounding of the true score may move; shards add up exactly; filtered <= raw; oracle ranks on a sample."""
    import os
    from KGE._b200.filter import build_filter
    rng = np.random.default_rng(23)
    ref, W, eng = make_pair(model, n_ent, R, d, dr, score=_score_for(model, score), loss=LOSSES["hinge"])
    pos = rand_triples(rng, 20000, n_ent, R)
    Xq = pos[:700].copy()
    Xq[::4] = rand_triples(rng, len(Xq[::4]), n_ent, R)
    Xq[Xq[:, 1] == 3, 1] = 4                                 # a relation no query asks for
    off, ids = build_filter(pos, Xq, side, n_ent, eng.device)
    s = 0 if side == "h" else 1
    t_raw = eng.rank_counts(Xq, s).cpu().numpy()
    t_flt = eng.rank_counts(Xq, s, off, ids).cpu().numpy()
    os.environ["KGE_RANK_GENERIC"] = "1"
    try:
        g_raw = eng.rank_counts(Xq, s).cpu().numpy()
        g_flt = eng.rank_counts(Xq, s, off, ids).cpu().numpy()
    finally:
        os.environ.pop("KGE_RANK_GENERIC")
    for a, b in ((t_raw, g_raw), (t_flt, g_flt)):
        diff = np.abs(a.astype(np.int64) - b)
        assert diff.max() <= 2 and (diff > 0).mean() <= 0.01, (diff.max(), (diff > 0).mean())
    assert (t_flt <= t_raw).all() and (t_flt >= 0).all()
    parts = sum(eng.rank_counts(Xq, s, off, ids, ent_lo=lo, ent_hi=hi).cpu().numpy()
                for lo, hi in [(0, n_ent // 3), (n_ent // 3, n_ent // 3 + 77), (n_ent // 3 + 77, n_ent)])
    assert np.array_equal(parts, t_flt)
    if model in ("TransR", "SE"):
        # the table rows come from a register-tiled GEMM whose every element is proj_fwd's own fmaf chain:
        # counts equal the warp-per-row table's bit for bit
        os.environ["KGE_PROJ_MATVEC"] = "1"
        try:
            assert np.array_equal(eng.rank_counts(Xq, s, off, ids).cpu().numpy(), t_flt)
        finally:
            os.environ.pop("KGE_PROJ_MATVEC")
    band = RankBand()
    for qi in range(40):
        oracle_rank_check(band, ref, W, Xq[qi], side, pos, int(t_flt[qi]) + 1, qi)
    print(model, score, side, band.summary())
    assert band.used <= 4, band.summary()


@pytest.mark.parametrize("model,d,dr", [("TransD", 50, 30), ("TransR", 40, 24), ("SE", 24, None)])
def test_projected_table_rank_in_relation_chunks(model, d, dr):
    """The projected tables of all relations rarely fit the workspace budget (256 MB; FB15k-237 TransR: 1.4 GB):
    relations are then taken a chunk at a time.  With the budget forced down to two relations per chunk
    (KGE_PROJ_BUDGET_KB) the counts must equal the one-chunk run exactly."""
    import os
    from KGE._b200.filter import build_filter
    n_ent = 900
    rng = np.random.default_rng(29)
    ref, W, eng = make_pair(model, n_ent, R, d, dr, score=_score_for(model, "l1"), loss=LOSSES["hinge"])
    pos = rand_triples(rng, 12000, n_ent, R)
    Xq = pos[:500]
    for side, s in (("h", 0), ("t", 1)):
        off, ids = build_filter(pos, Xq, side, n_ent, eng.device)
        one = eng.rank_counts(Xq, s, off, ids).cpu().numpy()
        dout = dr if (dr and model != "SE") else d
        per_rel_kb = (1024 * ((dout + 3) // 4 * 4) * 4) // 1024 + 1
        os.environ["KGE_PROJ_BUDGET_KB"] = str(2 * per_rel_kb)
        try:
            many = eng.rank_counts(Xq, s, off, ids).cpu().numpy()
        finally:
            os.environ.pop("KGE_PROJ_BUDGET_KB")
        assert np.array_equal(one, many), side
