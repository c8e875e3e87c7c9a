"""Row-lazy replay (kge_tables.lazy_t, kge_flush_rows): the training step that only touches the
looked-up entity rows must be BIT-identical to the dense step the reference's semantics call for
(start-of-batch constraint on the whole table + Keras Adam's dense decay / move of every row,
SURVEY.md A.3, App. B.2), on losses at every step and on w / m / v after the flush."""
import ctypes as C
import numpy as np
import pytest
import torch

from helpers import make_pair, rand_triples
from oracle import ref_ops as O

pytestmark = pytest.mark.gpu

CASES = [
    # model, score, loss, constraint, d, k, side
    ("TransE", O.LpDistance(1), O.PairwiseHingeLoss(1.0), True, 50, 2, "h+t"),
    ("TransE", O.LpDistance(2), O.PairwiseLogisticLoss(), False, 128, 4, "t"),
    ("DistMult", None, O.BinaryCrossEntropyLoss(), True, 200, 8, "h+t"),
    ("UM", O.LpDistancePow(2), O.PairwiseHingeLoss(1.0), True, 33, 3, "h"),
    ("TransD", O.LpDistancePow(2), O.PairwiseHingeLoss(1.0), True, 40, 4, "h+t"),
    ("TransH", O.LpDistance(1), O.PairwiseHingeLoss(1.0), False, 36, 2, "h+t"),
    ("RotatE", O.LpDistance(1), O.SelfAdversarialNegativeSamplingLoss(3.0, 1.0), True, 150, 6, "h+t"),
]


def _run(case, lazy, N=3000, R=7, B=48, steps=14, hot=True, opt=None):
    model, score, loss, con, d, k, side = case
    _, _, eng = make_pair(model, N, R, d, score=score, loss=loss, k=k, side=side, constraint=con, lazy=lazy)
    if opt == "sgd":                      # kge_cfg.optimizer (KGE.optimizers.SGD / Adagrad in the model API)
        eng.cfg.optimizer, eng.cfg.lr = 1, 0.05
    elif opt == "adagrad":
        eng.cfg.optimizer, eng.cfg.lr, eng.acc_init = 2, 0.05, 0.1
    rng = np.random.default_rng(5)
    losses = []
    for step in range(steps):
        X = rand_triples(rng, B, N, R)
        X[:, 0] = np.where(rng.random(B) < 0.5, X[:, 0] % 40, X[:, 0])     # a warm region and a cold tail
        if hot and step in (4, 9):
            X = np.concatenate([X, np.tile(X[:1], (70, 1))])                # one row with > 48 records
        if step == 6:                                                        # validation batch mid-way
            losses.append(float(eng.train_step(X, 11, step, is_train=False).item()))
            continue
        losses.append(float(eng.train_step(X, 11, step).item()))
    eng.flush()
    torch.cuda.synchronize()
    out = {}
    for s, t in eng.tab.items():
        out[s] = (t.storage.clone(), t.m.clone(), t.v.clone())
    return losses, out, eng


@pytest.mark.parametrize("case", CASES, ids=[c[0] + "-" + type(c[2]).__name__ + ("-con" if c[3] else "") for c in CASES])
def test_lazy_rows_equal_dense_step_bitwise(case):
    l_dense, t_dense, _ = _run(case, lazy=False)
    l_lazy, t_lazy, eng = _run(case, lazy=True)
    assert l_dense == l_lazy
    for s in t_dense:
        for name, a, b in zip("wmv", t_dense[s], t_lazy[s]):
            assert torch.equal(a, b), f"slot {s} {name}: {int((a != b).sum())} elements differ"
    last = eng.tab[0].last.cpu()
    assert int(last.min()) == int(last.max()) == eng.lazy_t == eng.adam_t       # everything flushed


def test_lazy_step_leaves_cold_rows_alone_until_flush():
    case = CASES[0]
    model, score, loss, con, d, k, side = case
    _, _, eng = make_pair(model, 5000, 5, d, score=score, loss=loss, k=1, side="t", constraint=con, lazy=True)
    w0 = eng.tab[0].storage.clone()
    rng = np.random.default_rng(2)
    X = rand_triples(rng, 16, 100, 5)                       # only entities < 100 (+ 16 random negatives)
    for step in range(3):
        eng.train_step(X, 3, step)
    torch.cuda.synchronize()
    last = eng.tab[0].last.cpu().numpy()
    touched = np.nonzero(last)[0]
    assert 0 < len(touched) <= 16 * 3 * 3
    cold = np.setdiff1d(np.arange(5000), touched)
    assert torch.equal(eng.tab[0].storage[cold], w0[cold])  # not even normalised yet
    assert eng._lazy_dirty
    nrm = eng.tab[0].view.norm(dim=1).cpu()                 # reading the table flushes it
    assert not eng._lazy_dirty
    assert float((nrm[cold] - 1).abs().max()) < 1e-6        # cold rows: the skipped normalisations were replayed
    assert int(eng.tab[0].last.min()) == 3


def test_lazy_rank_and_score_see_flushed_tables():
    case = CASES[2]
    outs = []
    for lazy in (False, True):
        eng = _run(case, lazy=lazy, steps=5, hot=False)[2]
        rng = np.random.default_rng(9)
        X = rand_triples(rng, 20, 3000, 7)
        eng.train_step(X, 1, 99)                            # leaves the lazy tables behind again
        sc = eng.score(X[:, 0], X[:, 1], X[:, 2]).cpu()
        cnt = eng.rank_counts(X, 1).cpu()
        outs.append((sc, cnt))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


def test_lazy_rejects_unsupported_and_out_of_order_calls():
    from KGE._b200 import lib as L
    with pytest.raises(RuntimeError, match="row-lazy"):
        _, _, eng = make_pair("TransR", 50, 3, 8, 8, score=O.LpDistancePow(2), loss=O.PairwiseHingeLoss(1.0), lazy=True)
        eng.train_step(rand_triples(np.random.default_rng(0), 8, 50, 3), 1, 0)
    with pytest.raises(RuntimeError, match="row-lazy"):
        _, _, eng = make_pair("TransH", 50, 3, 8, score=O.LpDistance(1), loss=O.PairwiseHingeLoss(1.0), constraint=True, lazy=True)
        eng.train_step(rand_triples(np.random.default_rng(0), 8, 50, 3), 1, 0)
    _, _, eng = make_pair("TransE", 50, 3, 8, score=O.LpDistance(1), loss=O.PairwiseHingeLoss(1.0), lazy=True)
    X = rand_triples(np.random.default_rng(0), 8, 50, 3)
    eng.train_step(X, 1, 0)
    eng.adam_t += 3                                         # skipping steps is refused
    with pytest.raises(RuntimeError, match="lazy_t"):
        eng.train_step(X, 1, 1)


def test_api_train_and_evaluate_identical_with_lazy_rows(tmp_path, monkeypatch):
    """KGE.models.* with KGE_B200_LAZY_ROWS=1: train() incl. validation / early stopping / checkpoints and
    evaluate() give the bits of the dense path."""
    from KGE.models.translating_based.TransE import TransE
    from toy_kg import train, val, metadata
    outs = []
    for flag in ("0", "1"):
        monkeypatch.setenv("KGE_B200_LAZY_ROWS", flag)
        m = TransE(embedding_params={"embedding_size": 12}, negative_ratio=2, corrupt_side="h+t")
        m.train(train, val, metadata, epochs=5, batch_size=4, early_stopping_rounds=10, seed=2, log_path=str(tmp_path / flag))
        assert m._engine.lazy == (flag == "1")
        ev = m.evaluate(val, "t", positive_X=np.concatenate([train, val]))
        outs.append(({k: v.numpy().copy() for k, v in m.model_weights.items()}, ev, list(m.val_loss_history)))
    for k in outs[0][0]:
        assert np.array_equal(outs[0][0][k], outs[1][0][k]), k
    assert outs[0][1] == outs[1][1] and outs[0][2] == outs[1][2]


@pytest.mark.parametrize("opt", ["sgd", "adagrad"])
@pytest.mark.parametrize("case", [CASES[0], CASES[2], CASES[4]], ids=["TransE", "DistMult", "TransD"])
def test_lazy_rows_equal_dense_step_bitwise_sgd_adagrad(case, opt):
    """SGD / Adagrad on row-lazy tables: a row without a gradient is not moved by these optimizers,
    so its skipped steps replay the start-of-batch constraint only — still bit-identical to the dense step."""
    l_dense, t_dense, _ = _run(case, lazy=False, opt=opt)
    l_lazy, t_lazy, eng = _run(case, lazy=True, opt=opt)
    assert l_dense == l_lazy
    for s in t_dense:
        for name, a, b in zip("wmv", t_dense[s], t_lazy[s]):
            assert torch.equal(a, b), f"slot {s} {name}: {int((a != b).sum())} elements differ"
