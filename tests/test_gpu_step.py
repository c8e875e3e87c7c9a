"""GPU parity of the fused training step / scoring against the CPU oracle
(tolerances from BASELINE.json north_star: scores and losses 1e-4 relative in fp32,
negatives bit-exact on the shared Philox stream)."""
import math
import numpy as np
import pytest
import torch

from helpers import make_pair, tables_of, rand_triples, rel_err, assert_weights_close
from oracle import ref_ops as O, ref_step as RS, ref_models as RM

pytestmark = pytest.mark.gpu

N, R = 300, 11
LOSSES = {
    "hinge": O.PairwiseHingeLoss(1.0), "logistic": O.PairwiseLogisticLoss(),
    "bce": O.BinaryCrossEntropyLoss(), "selfadv": O.SelfAdversarialNegativeSamplingLoss(3.0, 1.0),
    "square": O.SquareErrorLoss(),
}
SCORES = {"l1": O.LpDistance(1), "l2": O.LpDistance(2), "linf": O.LpDistance(math.inf),
          "l2pow": O.LpDistancePow(2), "l3": O.LpDistance(3), "dot": O.Dot()}
VECTOR_MODELS = ["UM", "TransE", "TransH", "TransD", "RotatE", "DistMult"]


def _score_for(model, name):
    return None if model in ("DistMult", "RESCAL") else SCORES[name]


@pytest.mark.parametrize("model", VECTOR_MODELS)
@pytest.mark.parametrize("score", ["l1", "l2", "linf", "l2pow", "l3", "dot"])
def test_score_hrt_matches_oracle(model, score):
    if model == "RotatE" and score == "dot":
        pytest.skip("complex dot is not a real score (tests/test_integration.py:47)")
    if model == "DistMult" and score != "dot":
        pytest.skip("DistMult has no score_fn")
    rng = np.random.default_rng(1)
    d = 50 if model != "RotatE" else 37
    ref, W, eng = make_pair(model, N, R, d, score=_score_for(model, score), loss=LOSSES["hinge"])
    X = rand_triples(rng, 257, N, R)
    got = eng.score(X[:, 0], X[:, 1], X[:, 2]).cpu()
    Xt = torch.as_tensor(X)
    with torch.no_grad():
        want = ref.score_hrt(W, Xt[:, 0], Xt[:, 1], Xt[:, 2])
    assert rel_err(got, want) < 1e-4


def _run_steps(model, score, loss, k, side, d, steps=4, B=64, constraint=True, d_rel=None, seed=11):
    rng = np.random.default_rng(3)
    ref, W, eng = make_pair(model, N, R, d, d_rel, score=_score_for(model, score), loss=LOSSES[loss], k=k,
                            side=side, constraint=constraint)
    opt = RS.TFAdam()
    for step in range(steps):
        X = rand_triples(rng, B, N, R)
        ids, k_eff = RS.sample_negatives(seed, step, X, k, side, N)
        l_ref, diag = RS.run_single_batch(ref, W, opt, X, ids, k_eff, side, True)
        l_gpu, neg = eng.train_step(X, seed, step, is_train=True, want_negatives=True)
        assert np.array_equal(neg.cpu().numpy(), ids), "negatives must be bit-exact"
        assert abs(float(l_gpu.item()) - l_ref) <= 1e-4 * max(1.0, abs(l_ref)), (step, float(l_gpu.item()), l_ref)
        got = tables_of(eng, model)
        for kk, v in W.data.items():
            assert_weights_close(got[kk], v, 1e-3, step + 1, (step, kk))
    return ref, W, eng


@pytest.mark.parametrize("model", VECTOR_MODELS)
def test_train_steps_default_config(model):
    score = {"UM": "l2pow", "TransE": "l2", "TransH": "l2pow", "TransD": "l2pow", "RotatE": "l1",
             "DistMult": "dot"}[model]
    loss = "selfadv" if model == "RotatE" else "hinge"
    _run_steps(model, score, loss, k=4, side="h+t", d=50 if model != "RotatE" else 21,
               constraint=(model != "TransH"))


@pytest.mark.parametrize("loss", list(LOSSES))
@pytest.mark.parametrize("side", ["h", "t", "h+t"])
def test_transe_losses_sides(loss, side):
    _run_steps("TransE", "l1", loss, k=6, side=side, d=50)


@pytest.mark.parametrize("score", ["l1", "l2", "linf", "l2pow", "l3", "dot"])
def test_transe_scores(score):
    _run_steps("TransE", score, "hinge", k=3, side="h+t", d=33)


@pytest.mark.parametrize("k,d", [(1, 50), (64, 200), (20, 300), (130, 16)])
def test_distmult_bce_shapes(k, d):
    _run_steps("DistMult", "dot", "bce", k=k, side="h+t" if k > 1 else "t", d=d, steps=3)


def test_transd_rectangular():
    _run_steps("TransD", "l2pow", "hinge", k=4, side="h+t", d=24, d_rel=40)
    _run_steps("TransD", "l2pow", "hinge", k=4, side="h+t", d=40, d_rel=24)


def test_rotate_wide_rows_selfadv():
    _run_steps("RotatE", "l1", "selfadv", k=64, side="h+t", d=500, steps=2, B=32)


def test_validation_batch_mutates_only_constraint():
    rng = np.random.default_rng(5)
    ref, W, eng = make_pair("TransE", N, R, 50, score=SCORES["l1"], loss=LOSSES["hinge"], k=2, side="t")
    X = rand_triples(rng, 32, N, R)
    ids, k_eff = RS.sample_negatives(9, 0, X, 2, "t", N)
    l_ref, _ = RS.run_single_batch(ref, W, RS.TFAdam(), X, ids, k_eff, "t", False)
    l_gpu = eng.train_step(X, 9, 0, is_train=False)
    assert abs(float(l_gpu.item()) - l_ref) <= 1e-4 * max(1.0, abs(l_ref))
    assert rel_err(tables_of(eng, "TransE")["ent_emb"], W.data["ent_emb"]) < 1e-6


def test_step_is_bitwise_deterministic():
    outs = []
    for _ in range(2):
        rng = np.random.default_rng(3)
        ref, W, eng = make_pair("DistMult", N, R, 200, loss=LOSSES["bce"], k=64, side="h+t")
        for step in range(3):
            eng.train_step(rand_triples(rng, 128, N, R), 5, step)
        outs.append(tables_of(eng, "DistMult"))
    for kk in outs[0]:
        assert torch.equal(outs[0][kk], outs[1][kk])


def test_negatives_override_is_used():
    rng = np.random.default_rng(8)
    ref, W, eng = make_pair("TransE", N, R, 50, score=SCORES["l1"], loss=LOSSES["hinge"], k=2, side="h+t")
    X = rand_triples(rng, 16, N, R)
    ids = rng.integers(0, N, 16 * 2)
    l_ref, _ = RS.run_single_batch(ref, W, RS.TFAdam(), X, ids, 2, "h+t", True)
    l_gpu, neg = eng.train_step(X, 0, 0, neg_override=ids, want_negatives=True)
    assert np.array_equal(neg.cpu().numpy(), ids)
    assert abs(float(l_gpu.item()) - l_ref) <= 1e-4 * max(1.0, abs(l_ref))


# ---- matrix-relation models and dense regularisers ----------------
MATRIX_DEFAULTS = {"SE": ("l1", "hinge"), "TransR": ("l2pow", "hinge"), "RESCAL": ("dot", "square")}


def _score_mat(model, name):
    return None if model == "RESCAL" else SCORES[name]


@pytest.mark.parametrize("model", list(MATRIX_DEFAULTS))
@pytest.mark.parametrize("constraint", [True, False])
def test_matrix_models_default_config(model, constraint):
    score, loss = MATRIX_DEFAULTS[model]
    rng = np.random.default_rng(13)
    ref, W, eng = make_pair(model, N, R, 24, 24, score=_score_mat(model, score), loss=LOSSES[loss], k=4, side="h+t",
                            constraint=constraint)
    opt = RS.TFAdam()
    for step in range(3):
        X = rand_triples(rng, 48, N, R)
        ids, k_eff = RS.sample_negatives(5, step, X, 4, "h+t", N)
        l_ref, _ = RS.run_single_batch(ref, W, opt, X, ids, k_eff, "h+t", True)
        l_gpu = eng.train_step(X, 5, step)
        assert abs(float(l_gpu.item()) - l_ref) <= 1e-4 * max(1.0, abs(l_ref)), (step, float(l_gpu.item()), l_ref)
        got = tables_of(eng, model)
        for kk, v in W.data.items():
            assert_weights_close(got[kk], v, 1e-3, step + 1, (step, kk))


@pytest.mark.parametrize("model,d,dr,k,side,loss", [
    ("TransR", 100, 100, 10, "h+t", "hinge"),      # C4 shape
    ("TransR", 40, 70, 3, "h", "logistic"),        # rectangular, clip active, unnormalised loss (clip_by_norm active)
    ("TransR", 70, 40, 5, "t", "selfadv"),
    ("SE", 50, 50, 6, "h+t", "bce"),
    ("SE", 33, 33, 2, "t", "hinge"),
    ("RESCAL", 64, 64, 8, "h+t", "square"),
    ("RESCAL", 30, 30, 3, "h", "logistic"),
    ("TransR", 200, 200, 2, "h+t", "hinge"),       # matrix too large to stage with Adam accumulators: global path
])
def test_matrix_models_shapes(model, d, dr, k, side, loss):
    score = {"TransR": "l2pow", "SE": "l1", "RESCAL": "dot"}[model]
    rng = np.random.default_rng(17)
    ref, W, eng = make_pair(model, N, R, d, dr, score=_score_mat(model, score), loss=LOSSES[loss], k=k, side=side)
    opt = RS.TFAdam()
    for step in range(2):
        X = rand_triples(rng, 40, N, R)
        ids, k_eff = RS.sample_negatives(21, step, X, k, side, N)
        l_ref, _ = RS.run_single_batch(ref, W, opt, X, ids, k_eff, side, True)
        l_gpu, neg = eng.train_step(X, 21, step, want_negatives=True)
        assert np.array_equal(neg.cpu().numpy(), ids)
        assert abs(float(l_gpu.item()) - l_ref) <= 1e-4 * max(1.0, abs(l_ref)), (step, float(l_gpu.item()), l_ref)
        got = tables_of(eng, model)
        for kk, v in W.data.items():
            assert_weights_close(got[kk], v, 1e-3, step + 1, (step, kk))


@pytest.mark.parametrize("model", list(MATRIX_DEFAULTS))
def test_matrix_score_hrt(model):
    score, loss = MATRIX_DEFAULTS[model]
    rng = np.random.default_rng(2)
    ref, W, eng = make_pair(model, N, R, 50, 30 if model == "TransR" else 50, score=_score_mat(model, score),
                            loss=LOSSES[loss])
    X = rand_triples(rng, 129, N, R)
    got = eng.score(X[:, 0], X[:, 1], X[:, 2]).cpu()
    Xt = torch.as_tensor(X)
    with torch.no_grad():
        want = ref.score_hrt(W, Xt[:, 0], Xt[:, 1], Xt[:, 2])
    assert rel_err(got, want) < 1e-4


@pytest.mark.parametrize("cw", [1.0, 0.25])
def test_transh_dense_regulariser(cw):
    """TransH constraint=True: soft constraints give dense gradients on all three tables and
    tf.clip_by_norm sees the densified tensors (TransH.py:202-211)."""
    rng = np.random.default_rng(23)
    ref, W, eng = make_pair("TransH", N, R, 50, score=SCORES["l2pow"], loss=LOSSES["hinge"], k=4, side="h+t",
                            constraint=True, cw=cw)
    opt = RS.TFAdam()
    for step in range(3):
        X = rand_triples(rng, 64, N, R)
        ids, k_eff = RS.sample_negatives(8, step, X, 4, "h+t", N)
        is_train = step != 1
        l_ref, _ = RS.run_single_batch(ref, W, opt, X, ids, k_eff, "h+t", is_train)
        l_gpu = eng.train_step(X, 8, step, is_train=is_train)
        assert abs(float(l_gpu.item()) - l_ref) <= 1e-4 * max(1.0, abs(l_ref)), (step, float(l_gpu.item()), l_ref)
        got = tables_of(eng, "TransH")
        for kk, v in W.data.items():
            assert_weights_close(got[kk], v, 1e-3, step + 1, (step, kk))


def test_matrix_step_is_bitwise_deterministic():
    outs = []
    for _ in range(2):
        rng = np.random.default_rng(3)
        ref, W, eng = make_pair("TransR", N, R, 48, 32, score=SCORES["l2pow"], loss=LOSSES["hinge"], k=6, side="h+t")
        for step in range(3):
            eng.train_step(rand_triples(rng, 96, N, R), 5, step)
        outs.append(tables_of(eng, "TransR"))
    for kk in outs[0]:
        assert torch.equal(outs[0][kk], outs[1][kk])


@pytest.mark.parametrize("model,B,k,n_rel", [
    ("DistMult", 512, 64, 11),        # long relation segments (> 32 ids): bitmap windows
    ("TransE", 2048, 40, 3),          # segments longer than one 32768-id window span
    ("TransE", 4096, 40, 2),          # hot rows (> 1024 records): ordered by the whole CTA, one 262144-id window
    ("UM", 8192, 40, 2),              # ... and over more than one window
    ("TransD", 64, 2, 11),            # short segments only: warp bitonic network
    ("TransR", 96, 6, 4),             # matrix gradients read the relation segments in order
])
def test_grouping_equals_stable_radix_sort(model, B, k, n_rel):
    """The counting sort (csrc/group_kernels.cuh) must hand the update kernels exactly the record
    order of cub's stable radix sort: weights after several steps are bit-identical."""
    from KGE._b200 import lib as L
    rng = np.random.default_rng(17)
    score = None if model == "DistMult" else SCORES["l2pow"]
    d, dr = (24, 20) if model == "TransR" else (40, None)
    outs = []
    # radix sort | four-kernel counting sort (default) | one-kernel cluster grouping (k_group_cluster, opt-in)
    for radix, cluster in ((1, 0), (0, 0), (0, 1)):
        _, _, eng = make_pair(model, 120, n_rel, d, dr, score=score, loss=LOSSES["hinge"], k=k, side="h+t")
        eng.set_option(L.OPT_RADIX_SORT, radix)
        eng.set_option(L.OPT_CLUSTER_GROUP, cluster)
        r2 = np.random.default_rng(17)
        try:
            for step in range(3):
                eng.train_step(rand_triples(r2, B, 120, n_rel), 5, step)
            torch.cuda.synchronize()
        finally:
            eng.set_option(L.OPT_RADIX_SORT, 0)
            eng.set_option(L.OPT_CLUSTER_GROUP, 0)
        outs.append({s: t.storage.clone() for s, t in eng.tab.items()})
    for o in outs[1:]:
        for s in outs[0]:
            assert torch.equal(outs[0][s], o[s]), f"slot {s}"


@pytest.mark.parametrize("model,score", [("DistMult", "dot"), ("TransE", "l1"), ("UM", "l2pow")])
def test_shadow_table_is_bit_identical_to_constraint_pass(model, score):
    """kge_tables.wn (normalised rows written by the row update for the next batch) must change
    nothing: train / validation / external assign sequences give the same bits with the shadow and
    with the full-table constraint pass every batch (KGE_OPT_NO_SHADOW)."""
    from KGE._b200 import lib as L
    outs, losses = [], []
    for no_shadow in (1, 0):
        rng = np.random.default_rng(23)
        _, _, eng = make_pair(model, N, R, 40, score=_score_for(model, score), loss=LOSSES["hinge"], k=6, side="h+t")
        assert eng.tab[0].wn is not None
        eng.set_option(L.OPT_NO_SHADOW, no_shadow)
        ls = []
        try:
            for step in range(8):
                X = rand_triples(rng, 96, N, R)
                if step == 3:                       # validation batch: assigns the constraint, no update
                    ls.append(float(eng.train_step(X, 5, step, is_train=False).item()))
                    continue
                if step == 5:                       # caller overwrites weights: the shadow must be dropped
                    eng.tab[0].view.mul_(1.5)
                    eng.tab[0].touch()
                ls.append(float(eng.train_step(X, 5, step).item()))
                if not no_shadow:
                    assert eng.tab[0].wn_valid
            torch.cuda.synchronize()
        finally:
            eng.set_option(L.OPT_NO_SHADOW, 0)
        outs.append({s: t.storage.clone() for s, t in eng.tab.items()})
        losses.append(ls)
    assert losses[0] == losses[1]
    for s in outs[0]:
        assert torch.equal(outs[0][s], outs[1][s]), f"slot {s}"


@pytest.mark.parametrize("model,d,dr,k,side,loss", [
    ("TransR", 100, 100, 10, "h+t", "hinge"), ("TransR", 40, 70, 3, "h", "logistic"), ("TransR", 70, 40, 5, "t", "selfadv"),
    ("TransR", 50, 30, 33, "h+t", "selfadv"), ("SE", 50, 50, 6, "h+t", "bce"), ("SE", 33, 33, 2, "t", "hinge"),
    ("SE", 24, 24, 9, "h", "logistic"), ("TransR", 160, 130, 4, "h+t", "hinge"),
])
def test_projection_step_kernel_equals_per_negative_kernel(model, d, dr, k, side, loss):
    """k_step_proj (four vectors per warp through every mat-vec, csrc/step_proj_kernels.cuh) against k_step_mat
    (KGE_STEP_MAT_OLD=1): every projection is the same fmaf chain, only the fixed order of the sums over a
    positive's negatives differs — losses to rounding, weights at the parity tolerance, validation batches too."""
    import os
    score = {"TransR": "l2pow", "SE": "l1"}[model]
    outs = []
    for old in (False, True):
        if old:
            os.environ["KGE_STEP_MAT_OLD"] = "1"
        try:
            rng = np.random.default_rng(29)
            ref, W, eng = make_pair(model, N, R, d, dr, score=_score_mat(model, score), loss=LOSSES[loss], k=k, side=side)
            losses = []
            for step in range(3):
                X = rand_triples(rng, 72, N, R)
                l, neg = eng.train_step(X, 4, step, is_train=step != 1, want_negatives=True)
                losses.append((float(l.item()), neg.cpu().numpy().copy()))
            outs.append((losses, tables_of(eng, model)))
        finally:
            os.environ.pop("KGE_STEP_MAT_OLD", None)
    for (la, na), (lb, nb) in zip(outs[0][0], outs[1][0]):
        assert np.array_equal(na, nb)
        assert abs(la - lb) <= 2e-6 * max(1.0, abs(lb)), (la, lb)
    for kk in outs[0][1]:
        assert_weights_close(outs[0][1][kk], outs[1][1][kk], 1e-3, 2, kk)


def test_transr_at_the_reference_default_init():
    """The reference initialises rel_proj to the rectangular identity (TransR.py:125-128), so after the
    start-of-batch normalisation ||e M_r|| sits ON the clip threshold of TransR.py:187-189 and rounding decides
    the branch (helpers.make_pair perturbs the matrices for that reason).  At the unperturbed init the clip is
    still continuous in VALUE: losses agree to 1e-4 over three steps and every weight stays inside Adam's
    per-step bound; how many elements leave the 1e-4 band is printed, not asserted (the gradient of the clip
    jumps where the branch flips)."""
    from KGE._b200 import engine as E, spec
    rng = np.random.default_rng(31)
    d, dr, k = 40, 24, 4
    ref = RM.MODELS["TransR"](N, R, d, dr, score_fn=SCORES["l2pow"], loss_fn=LOSSES["hinge"], constraint=True,
                              constraint_weight=1.0)
    W = ref.init_weights(11)
    eye = W.data["rel_proj"].reshape(R, d, dr)
    assert torch.equal(eye[0], torch.eye(d, dr)), "rel_proj must start at the rectangular identity"
    cfg = spec.make_cfg("TransR", N, R, d, dr, score_fn=SCORES["l2pow"], loss_fn=LOSSES["hinge"], negative_ratio=k,
                        corrupt_side="h+t", constraint=True, constraint_weight=1.0, rotate_limit=ref.limit)
    eng = E.Engine(cfg, {spec.SLOTS["TransR"][kk]: v for kk, v in W.data.items()})
    opt = RS.TFAdam()
    for step in range(3):
        X = rand_triples(rng, 64, N, R)
        ids, k_eff = RS.sample_negatives(9, step, X, k, "h+t", N)
        l_ref, _ = RS.run_single_batch(ref, W, opt, X, ids, k_eff, "h+t", True)
        l_gpu = float(eng.train_step(X, 9, step).item())
        assert abs(l_gpu - l_ref) <= 1e-4 * max(1.0, abs(l_ref)), (step, l_gpu, l_ref)
        got = tables_of(eng, "TransR")
        for kk, v in W.data.items():
            diff = (got[kk].double() - torch.as_tensor(v).double()).abs()
            scale = float(torch.as_tensor(v).abs().max()) + 1e-12
            print("TransR identity init, step %d, %s: max|dW| %.3g, outside 1e-4: %.5f" % (
                step, kk, float(diff.max()), float((diff > 1e-4 * scale).double().mean())))
            assert float(diff.max()) <= (step + 1) * 1e-3 * 1.01 + 1e-6 * scale, (step, kk, float(diff.max()))


@pytest.mark.parametrize("d,k,side,loss,B", [
    (300, 64, "h+t", "selfadv", 40), (300, 6, "t", "hinge", 72), (130, 64, "h+t", "bce", 33), (130, 5, "h", "logistic", 64),
    (500, 16, "h+t", "selfadv", 24), (140, 1, "t", "hinge", 50),
])
def test_rotate_streamed_kernel_equals_register_kernel(d, k, side, loss, B):
    """k_step_rot (csrc/step_rot_kernels.cuh: rows streamed chunk by chunk, accumulators in shared memory) against
    k_step's register-resident RotatE path (KGE_STEP_ROT_OLD=1): the same expressions in the same order — losses and
    negatives bitwise equal over train / validation / train steps, tables to rounding — and both against the oracle."""
    import os
    outs = []
    for old in (False, True):
        if old:
            os.environ["KGE_STEP_ROT_OLD"] = "1"
        try:
            rng = np.random.default_rng(37)
            ref, W, eng = make_pair("RotatE", N, R, d, None, score=SCORES["l1"], loss=LOSSES[loss], k=k, side=side)
            opt = RS.TFAdam()
            losses = []
            for step in range(3):
                X = rand_triples(rng, B, N, R)
                l, neg = eng.train_step(X, 6, step, is_train=step != 1, want_negatives=True)
                losses.append((float(l.item()), neg.cpu().numpy().copy()))
                if not old:
                    ids, k_eff = RS.sample_negatives(6, step, X, k, side, N)
                    l_ref, _ = RS.run_single_batch(ref, W, opt, X, ids, k_eff, side, step != 1)
                    assert np.array_equal(neg.cpu().numpy(), ids)
                    assert abs(losses[-1][0] - l_ref) <= 1e-4 * max(1.0, abs(l_ref)), (step, losses[-1][0], l_ref)
            got = tables_of(eng, "RotatE")
            if not old:
                for kk, v in W.data.items():
                    assert_weights_close(got[kk], v, 1e-3, 2, kk)
            outs.append((losses, got))
        finally:
            os.environ.pop("KGE_STEP_ROT_OLD", None)
    for (la, na), (lb, nb) in zip(outs[0][0], outs[1][0]):
        assert np.array_equal(na, nb)
        assert abs(la - lb) <= 1e-6 * max(1.0, abs(lb)), (la, lb)
    for kk in outs[0][1]:          # same expressions; the compiler contracts a few multiply-adds differently in the two kernels
        assert_weights_close(outs[0][1][kk], outs[1][1][kk], 1e-3, 2, "streamed vs register kernel, " + kk)
        assert float((outs[0][1][kk] - outs[1][1][kk]).abs().max()) <= 2e-5


@pytest.mark.parametrize("model,d,k,side,score,loss,B", [
    ("TransE", 50, 1, "t", "l1", "hinge", 4096), ("TransE", 50, 1, "h", "l2", "logistic", 333), ("TransE", 64, 2, "t", "l2pow", "bce", 77),
    ("UM", 33, 1, "h", "l2pow", "hinge", 500), ("TransE", 13, 2, "t", "linf", "square", 129), ("UM", 50, 1, "t", "l3", "hinge", 5000),
])
def test_two_positives_per_warp_kernel_equals_k_step_bitwise(model, d, k, side, score, loss, B):
    """k_step_narrow (csrc/step_narrow_kernels.cuh: rows of <= 64 floats, one 16-lane half per positive) against
    k_step (KGE_STEP_NARROW_OFF=1): same loads, expressions and butterfly sums, the in-kernel tail in k_step's order
    (also B > 4096: per-positive partials + k_reduce) — losses, negatives and tables bitwise equal over train /
    validation / train steps; the new kernel also against the oracle."""
    import os
    outs = []
    for off in (False, True):
        if off:
            os.environ["KGE_STEP_NARROW_OFF"] = "1"
        try:
            rng = np.random.default_rng(43)
            ref, W, eng = make_pair(model, N, R, d, None, score=SCORES[score], loss=LOSSES[loss], k=k, side=side)
            opt = RS.TFAdam()
            losses = []
            for step in range(3):
                X = rand_triples(rng, B, N, R)
                l, neg = eng.train_step(X, 8, step, is_train=step != 1, want_negatives=True)
                losses.append((float(l.item()), neg.cpu().numpy().copy()))
                if not off and B <= 600:
                    ids, k_eff = RS.sample_negatives(8, step, X, k, side, N)
                    l_ref, _ = RS.run_single_batch(ref, W, opt, X, ids, k_eff, side, step != 1)
                    assert np.array_equal(neg.cpu().numpy(), ids)
                    assert abs(losses[-1][0] - l_ref) <= 1e-4 * max(1.0, abs(l_ref)), (step, losses[-1][0], l_ref)
            outs.append((losses, tables_of(eng, model)))
        finally:
            os.environ.pop("KGE_STEP_NARROW_OFF", None)
    for (la, na), (lb, nb) in zip(outs[0][0], outs[1][0]):
        assert np.array_equal(na, nb) and la == lb, (la, lb)
    for kk in outs[0][1]:
        assert torch.equal(outs[0][1][kk], outs[1][1][kk]), kk


@pytest.mark.parametrize("model,d,k,side,score,loss,B,n_ent", [
    ("TransE", 50, 1, "t", "l1", "hinge", 4096, N), ("TransE", 50, 8, "h+t", "l2", "logistic", 600, 61),
    ("UM", 33, 3, "h", "l2pow", "hinge", 500, N), ("DistMult", 13, 2, "t", None, "bce", 129, 40), ("TransH", 24, 4, "h+t", "l1", "hinge", 200, N),
])
def test_two_rows_per_warp_update_equals_k_update_fast(model, d, k, side, score, loss, B, n_ent):
    """k_update_narrow (rows of <= 64 floats, one 16-lane half per table row) against k_update_fast<1, 0>
    (KGE_UPDATE_NARROW_OFF=1): records added in the same order, same Adam, same shadow rows — tables and Adam slots
    bitwise equal; a 61 / 40-entity graph makes hot rows (k_update_heavy's) and rows without records share warps."""
    import os
    outs = []
    for off in (False, True):
        if off:
            os.environ["KGE_UPDATE_NARROW_OFF"] = "1"
        try:
            rng = np.random.default_rng(47)
            ref, W, eng = make_pair(model, n_ent, R, d, None, score=_score_for(model, score) if score else None, loss=LOSSES[loss], k=k, side=side)
            losses = [float(eng.train_step(rand_triples(rng, B, n_ent, R), 8, step, is_train=step != 1).item()) for step in range(4)]
            eng.flush()
            outs.append((losses, tables_of(eng, model), {s: (t.m.clone(), t.v.clone()) for s, t in eng.tab.items() if t.m is not None}))
        finally:
            os.environ.pop("KGE_UPDATE_NARROW_OFF", None)
    assert outs[0][0] == outs[1][0]
    for kk in outs[0][1]:
        assert torch.equal(outs[0][1][kk], outs[1][1][kk]), kk
    for s in outs[0][2]:
        assert torch.equal(outs[0][2][s][0], outs[1][2][s][0]) and torch.equal(outs[0][2][s][1], outs[1][2][s][1]), s
