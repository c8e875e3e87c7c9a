"""The reference's own test assertions, run against the CPU oracle:
tests/test_score.py, test_loss.py, test_constraint.py, test_metrics.py, test_model.py."""
import math
import numpy as np
import pytest
import torch
from oracle import ref_ops as O, ref_models as RM, ref_step as RS

torch.manual_seed(0)


@pytest.mark.parametrize("p", [1, 2, math.inf])
def test_lp_scores_nonpositive_finite(p):              # tests/test_score.py:7-38
    x, y = torch.rand(10, 5) * 2 - 1, torch.rand(10, 5) * 2 - 1
    for fn in (O.LpDistance(p), O.LpDistancePow(p)):
        s = fn(x, y)
        assert s.shape == (10,) and bool((s <= 0).all()) and bool(torch.isfinite(s).all())
        sc = fn((x, x), (y, y))                                # complex inputs
        assert sc.shape == (10,) and bool((sc <= 0).all())


def test_dot_finite():                                 # tests/test_score.py:40-49
    s = O.Dot()(torch.rand(10, 5), torch.rand(10, 5))
    assert s.shape == (10,) and bool(torch.isfinite(s).all())


@pytest.mark.parametrize("loss", [O.PairwiseHingeLoss(1.0), O.PairwiseLogisticLoss(), O.BinaryCrossEntropyLoss(),
                                  O.SelfAdversarialNegativeSamplingLoss(3.0, 1.0), O.SquareErrorLoss()])
def test_losses_scalar_nonnegative(loss):              # tests/test_loss.py:8-70
    v = loss(torch.rand(10), torch.rand(20))
    assert v.dim() == 0 and math.isfinite(float(v)) and float(v) >= 0


def test_loss_normalisations():
    pos, neg = torch.tensor([0.5, -0.2]), torch.tensor([0.1, 0.9, -0.3, 0.4])
    a = float(O.PairwiseHingeLoss(1.0)(pos, neg))
    want = (max(0, 1 + .1 - .5) + max(0, 1 + .9 - .5) + max(0, 1 - .3 + .2) + max(0, 1 + .4 + .2)) / 4
    assert abs(a - want) < 1e-6


@pytest.mark.parametrize("p", [1, 2, math.inf])
def test_constraints(p):                               # tests/test_constraint.py:8-66
    X = torch.rand(10, 5)
    Y = O.normalized_embeddings(X, p, 1, -1)
    assert Y.shape == X.shape and float((O._norm(Y, p, -1) - 1).abs().max()) <= 1e-6
    assert float(O.soft_constraint(X, p, 1, -1)) >= 0
    assert float((O._norm(O.clip_constraint(X * 3, p, 1, -1), p, -1) - 1).max()) <= 1e-6
    r = O.Lp_regularization(X, 2, -1)
    assert r.shape == (10,) and bool((r >= 0).all())


def test_metrics_ranges():                             # tests/test_metrics.py:7-51
    m = RS.metrics(np.random.default_rng(0).integers(1, 1000, 10))
    assert 0 < m["mean_reciprocal_rank"] <= 1 and m["mean_rank"] >= 1 and m["median_rank"] >= 1
    assert m["geometric_mean_rank"] >= 1 and m["harmonic_mean_rank"] >= 1 and m["std_rank"] >= 0
    assert all(m[k] <= 1 for k in ("hit@1", "hit@3", "hit@10"))
    assert set(m) == {"mean_rank", "mean_reciprocal_rank", "median_rank", "geometric_mean_rank",
                      "harmonic_mean_rank", "std_rank", "hit@1", "hit@3", "hit@10"}


@pytest.mark.parametrize("name", list(RM.MODELS))
def test_models_score_shapes_and_filtered_rank(name):  # tests/test_model.py:19-283
    N, R, d = 12, 11, 16
    sf = None if name in ("DistMult", "RESCAL") else O.LpDistance(2)
    m = RM.MODELS[name](N, R, d, d, score_fn=sf, loss_fn=O.PairwiseHingeLoss(1.0))
    W = m.init_weights(3)
    assert set(W.data) == set(m.shapes())
    rng = np.random.default_rng(1)
    X = np.stack([rng.integers(0, N, 12), rng.integers(0, R, 12), rng.integers(0, N, 12)], 1)
    Xt = torch.as_tensor(X)
    with torch.no_grad():
        s = m.score_hrt(W, Xt[:3, 0], Xt[:3, 1], Xt[:3, 2])
    assert s.shape == (3,) and bool(torch.isfinite(s).all())
    assert RS.get_scores_all(m, W, X[0], "h").shape == (N,)
    opt = RS.TFAdam()
    ids, k = RS.sample_negatives(1, 0, X, 2, "h+t", N)
    loss, _ = RS.run_single_batch(m, W, opt, X, ids, k, "h+t", True)
    assert math.isfinite(loss)
    for side in ("h", "t"):
        raw, flt = RS.get_rank(m, W, X[0], None, side), RS.get_rank(m, W, X[0], X, side)
        assert 1 <= flt <= raw


def test_tf_adam_sparse_moves_untouched_rows():
    opt = RS.TFAdam()
    w = torch.ones(4, 2)
    for _ in range(2):
        opt.begin_step()
        opt.apply_sparse("w", w, torch.tensor([1, 1]), torch.ones(2, 2))
    opt.begin_step()
    before = w.clone()
    opt.apply_sparse("w", w, torch.tensor([3]), torch.ones(1, 2))
    assert not torch.equal(w[1], before[1])     # row 1 untouched this step, still drifts (m != 0)
    assert torch.equal(w[0], before[0])         # never touched: m == 0, stays
