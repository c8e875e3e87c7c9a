"""Optimizers other than Adam (BaseModel.py:243-246 accepts any Keras optimizer): SGD and
Adagrad row updates of the CUDA step against the oracle's TF 2.5 sparse-apply restatement."""
import numpy as np
import pytest
import torch

from helpers import make_pair, tables_of, rand_triples, rel_err
from oracle import ref_ops as O, ref_step as RS

pytestmark = pytest.mark.gpu

N, R = 300, 11
CASES = [
    ("TransE", O.LpDistance(1), O.PairwiseHingeLoss(1.0), 40, None, True),
    ("DistMult", None, O.BinaryCrossEntropyLoss(), 40, None, True),
    ("TransR", O.LpDistancePow(2), O.PairwiseHingeLoss(1.0), 12, 10, True),          # matrix table update
    ("TransH", O.LpDistance(2), O.PairwiseLogisticLoss(), 24, None, True),            # dense regulariser passes
    ("RotatE", O.LpDistance(1), O.SelfAdversarialNegativeSamplingLoss(3.0, 1.0), 30, None, True),
]


@pytest.mark.parametrize("optim", ["sgd", "adagrad"])
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_step_matches_oracle(case, optim):
    from KGE._b200 import lib as L
    model, score, loss, d, dr, con = case
    ref, W, eng = make_pair(model, N, R, d, dr, score=score, loss=loss, k=4, side="h+t", constraint=con)
    if optim == "sgd":
        opt = RS.TFSGD(lr=0.05)
        eng.cfg.optimizer, eng.cfg.lr = L.OPTIM_SGD, 0.05
    else:
        opt = RS.TFAdagrad(lr=0.05, initial_accumulator_value=0.1, eps=1e-7)
        eng.cfg.optimizer, eng.cfg.lr, eng.cfg.eps, eng.acc_init = L.OPTIM_ADAGRAD, 0.05, 1e-7, 0.1
    rng = np.random.default_rng(11)
    for step in range(4):
        X = rand_triples(rng, 64, N, R)
        X[:8] = X[0]                                          # duplicated rows: de-duplicated sum
        ids, k_eff = RS.sample_negatives(9, step, X, 4, "h+t", N)
        l_ref, _ = RS.run_single_batch(ref, W, opt, X, ids, k_eff, "h+t", True)
        l_gpu = eng.train_step(X, 9, step)
        assert abs(float(l_gpu.item()) - l_ref) <= 1e-4 * max(1.0, abs(l_ref)), (step, float(l_gpu.item()), l_ref)
    got = tables_of(eng, model)
    for kk, v in W.data.items():
        assert rel_err(got[kk], v) < 2e-5, (kk, rel_err(got[kk], v))
    if optim == "adagrad":                                    # the accumulator slot too
        from KGE._b200 import spec
        for kk, s in spec.SLOTS[model].items():
            acc = eng.tab[s].v[:, :eng.tab[s].cols].reshape(W.data[kk].shape).cpu()
            assert rel_err(acc, opt.acc[kk]) < 2e-5, kk


def test_api_accepts_optimizer_objects_and_names(tmp_path):
    from KGE import optimizers
    from KGE.models.translating_based.TransE import TransE
    from toy_kg import train, val, metadata
    for opt in (optimizers.SGD(0.05), optimizers.Adagrad(0.05), "SGD", "Adagrad", optimizers.Adam(1e-3)):
        m = TransE(embedding_params={"embedding_size": 8}, negative_ratio=2, corrupt_side="h+t")
        m.train(train, val, metadata, epochs=2, batch_size=4, optimizer=opt, seed=1, log_path=str(tmp_path))
        assert len(m.val_loss_history) >= 1 and np.isfinite(m.val_loss_history[-1])
    with pytest.raises(NotImplementedError):
        optimizers.adam_fields(optimizers.SGD(0.1, momentum=0.9))
    with pytest.raises(NotImplementedError):
        optimizers.adam_fields("RMSprop")


def test_lazy_rows_take_other_optimizers():
    """Row-lazy tables cover SGD / Adagrad too (bitwise against the dense step: tests/test_gpu_lazy.py)."""
    from KGE._b200 import lib as L
    _, _, eng = make_pair("TransE", 50, 3, 8, score=O.LpDistance(1), loss=O.PairwiseHingeLoss(1.0), lazy=True)
    eng.cfg.optimizer = L.OPTIM_SGD
    assert np.isfinite(float(eng.train_step(rand_triples(np.random.default_rng(0), 8, 50, 3), 1, 0).item()))
