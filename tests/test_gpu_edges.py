"""Edge cases of the batch hot path: the smallest and most ragged batches, the smallest graph, query lists of one,
filters that remove every candidate, and what the C-ABI refuses (empty batches / query lists: the reference's
tf.data pipeline never yields an empty batch, BaseModel.py:133-145, and its get_rank takes one triple)."""
import numpy as np
import pytest
import torch

from helpers import make_pair, rand_triples, RankBand, oracle_rank_check, assert_weights_close, tables_of
from oracle import ref_ops as O, ref_step as RS
from test_gpu_step import SCORES, LOSSES, _score_for

pytestmark = pytest.mark.gpu


def _steps(model, score, loss, N, R, d, B_list, k, side, d_rel=None, seed=5):
    rng = np.random.default_rng(17)
    ref, W, eng = make_pair(model, N, R, d, d_rel, score=_score_for(model, score), loss=LOSSES[loss], k=k, side=side)
    opt = RS.TFAdam()
    for step, B in enumerate(B_list):
        X = rand_triples(rng, B, N, R)
        ids, k_eff = RS.sample_negatives(seed, step, X, k, side, N)
        l_ref, _ = RS.run_single_batch(ref, W, opt, X, ids, k_eff, side, True)
        l_gpu, neg = eng.train_step(X, seed, step, is_train=True, want_negatives=True)
        assert np.array_equal(neg.cpu().numpy(), ids), ("negatives must be bit-exact", model, B)
        assert abs(float(l_gpu.item()) - l_ref) <= 1e-4 * max(1.0, abs(l_ref)), (model, B, float(l_gpu.item()), l_ref)
        got = tables_of(eng, model)
        # batch sizes that are not powers of two make the per-record gradient magnitude c / B inexact, so sums of
        # equal-magnitude L1 gradients no longer cancel to exactly 0 in every summation order, and Adam turns the
        # 1e-8 residue into a visible fraction of lr (helpers.assert_weights_close); the relation tables here have
        # 250 elements or fewer, 1 % would be two of them: 5 % outside 1e-4, the max bound unchanged
        for kk, v in W.data.items():
            assert_weights_close(got[kk], v, 1e-3, step + 1, (model, B, kk), max_frac=0.05)


@pytest.mark.parametrize("model,score,loss,d,d_rel", [
    ("TransE", "l1", "hinge", 50, None), ("UM", "l2pow", "hinge", 50, None), ("DistMult", "dot", "bce", 200, None),
    ("RotatE", "l1", "selfadv", 21, None), ("TransH", "l2pow", "hinge", 24, None), ("TransD", "l2pow", "hinge", 20, 12),
    ("TransR", "l2pow", "hinge", 16, 12), ("SE", "l1", "hinge", 16, None), ("RESCAL", "dot", "square", 16, None)])
def test_one_positive_one_negative(model, score, loss, d, d_rel):
    """B = 1, one negative on one side: every kernel with a single live positive, three steps."""
    _steps(model, score, loss, 60, 3, d, [1, 1, 1], k=1, side="t", d_rel=d_rel)


@pytest.mark.parametrize("model,score,loss,d,k,side", [
    ("TransE", "l1", "hinge", 50, 1, "h"),          # the two-positives-per-warp kernel (16 per CTA)
    ("TransE", "l2", "logistic", 100, 3, "h+t"),    # k_step, 8 positives per CTA
    ("DistMult", "dot", "bce", 200, 5, "h+t"),
    ("RotatE", "l1", "selfadv", 64, 6, "h+t")])
def test_ragged_batches_around_cta_boundaries(model, score, loss, d, k, side):
    """A different batch size every step (the reference's last batch of an epoch is ragged, data_utils.py:176-196):
    one below / at / one above the positives a CTA takes, and sizes that leave one warp-half idle."""
    _steps(model, score, loss, 120, 5, d, [7, 8, 9, 15, 16, 17, 33, 2], k=k, side=side)


@pytest.mark.parametrize("model,score", [("TransE", "l1"), ("DistMult", "dot"), ("TransR", "l2pow")])
def test_smallest_graph(model, score):
    """Two entities, one relation: every negative is one of the two entities, every row is touched by every record."""
    loss = "bce" if model == "DistMult" else "hinge"
    _steps(model, score, loss, 2, 1, 16, [4, 3], k=2, side="h+t", d_rel=8 if model == "TransR" else None)
    ref, W, eng = make_pair(model, 2, 1, 16, 8 if model == "TransR" else None, score=_score_for(model, score), loss=LOSSES[loss])
    X = np.array([[0, 0, 1], [1, 0, 1], [0, 0, 0]], dtype=np.int64)
    band = RankBand()
    for side, s in (("h", 0), ("t", 1)):
        cnt = eng.rank_counts(X, s).cpu().numpy()
        assert ((cnt >= 0) & (cnt <= 1)).all()
        for q, x in enumerate(X):
            oracle_rank_check(band, ref, W, x, side, None, int(cnt[q]) + 1, (model, side, q))
    print(band.summary())


@pytest.mark.parametrize("model,score,d", [("TransE", "l1", 50), ("DistMult", "dot", 64), ("RotatE", "l2", 20), ("TransH", "l2pow", 24),
                                           ("TransR", "l2pow", 16)])
def test_single_query_and_fully_filtered_query(model, score, d):
    """Q = 1 on every ranking path; a query whose filter lists EVERY entity ranks first (count 0) whatever the
    scores; a query with an empty filter equals the raw rank."""
    from KGE._b200.filter import build_filter
    N, R = 131, 4
    rng = np.random.default_rng(23)
    ref, W, eng = make_pair(model, N, R, d, score=_score_for(model, score), loss=LOSSES["hinge"])
    x = np.array([[5, 2, 77]], dtype=np.int64)
    band = RankBand()
    for side, s, fs, cs in (("h", 0, 2, 0), ("t", 1, 0, 2)):
        raw = eng.rank_counts(x, s).cpu().numpy()
        assert raw.shape == (1,)
        oracle_rank_check(band, ref, W, x[0], side, None, int(raw[0]) + 1, (model, side, "raw"))
        every = np.zeros((N, 3), dtype=np.int64)
        every[:, 1], every[:, fs], every[:, cs] = x[0, 1], x[0, fs], np.arange(N)
        off, ids = build_filter(every, x, side, N, eng.device)
        assert int(eng.rank_counts(x, s, off, ids).item()) == 0
        other = every.copy()
        other[:, 1] = (x[0, 1] + 1) % R                     # same entities under another relation: nothing filtered
        off, ids = build_filter(other, x, side, N, eng.device)
        assert int(eng.rank_counts(x, s, off, ids).item()) == int(raw[0])
        some = every[rng.permutation(N)[: N // 2]]
        off, ids = build_filter(some, x, side, N, eng.device)
        got = int(eng.rank_counts(x, s, off, ids).item())
        oracle_rank_check(band, ref, W, x[0], side, some, got + 1, (model, side, "half"))
    print(band.summary())


def test_empty_batches_and_query_lists_are_refused():
    _, _, eng = make_pair("TransE", 40, 3, 16, score=SCORES["l1"], loss=LOSSES["hinge"])
    with pytest.raises(ValueError, match="empty batch"):          # host mirror, before the step counter moves
        eng.train_step(np.zeros((0, 3), np.int64), 1, 0)
    with pytest.raises(ValueError, match="empty batch"):
        eng.train_step_host(torch.zeros((0, 3), dtype=torch.int64).pin_memory(), 1, 0, torch.zeros(1).pin_memory())
    with pytest.raises(RuntimeError, match="invalid"):
        eng.rank_counts(np.zeros((0, 3), np.int64), 0)
    with pytest.raises(RuntimeError, match="entity shard"):
        eng.rank_counts(np.array([[1, 1, 1]], np.int64), 0, ent_lo=10, ent_hi=10)
    # and the state is untouched by the refused calls: the next step equals a fresh engine's first step
    _, _, fresh = make_pair("TransE", 40, 3, 16, score=SCORES["l1"], loss=LOSSES["hinge"])
    X = rand_triples(np.random.default_rng(1), 9, 40, 3)
    a = eng.train_step(X, 1, 0)
    b = fresh.train_step(X, 1, 0)
    assert torch.equal(a, b)
    for kk, v in tables_of(fresh, "TransE").items():
        assert torch.equal(tables_of(eng, "TransE")[kk], v)


def test_evaluate_on_an_empty_split_returns_no_ranks(tmp_path):
    """KGE API: the reference's evaluate loops over eval_X (BaseModel.py:602-604); an empty split gives an empty
    rank list (and NaN metrics from numpy), not an error from the library."""
    import warnings
    from KGE.models.translating_based.TransE import TransE
    from toy_kg import train, val, metadata
    m = TransE(embedding_params={"embedding_size": 8}, negative_ratio=2, corrupt_side="h+t")
    m.train(train_X=train, val_X=val, metadata=metadata, epochs=1, batch_size=4, log_path=str(tmp_path))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        out = m.evaluate(np.zeros((0, 3), np.int64), "t", positive_X=train)
    assert len(m.last_ranks) == 0 and set(out) >= {"mean_rank", "hit@10"}
