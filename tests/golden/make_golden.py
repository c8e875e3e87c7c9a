#!/usr/bin/env python
"""Generate tests/golden/*.npz by executing the REFERENCE'S OWN PYTHON SOURCE
(/root/reference/KGE, unmodified, imported from where it lies) on fixed seeded inputs.

TensorFlow 2.5 cannot be installed here, so `import tensorflow` inside the reference
resolves to oracle/tf_shim/tensorflow (torch-backed, TF semantics of SURVEY.md App. B).
The fixtures therefore pin the reference's *formulas, op order, normalisations, row
layout of negatives, step sequencing and rank rule* — not TensorFlow's kernels.

Run from the repo root in the build container (needs /root/reference):
    python tests/golden/make_golden.py
Every fixture stores its inputs explicitly (weights, triples, negatives), so the tests
need neither the reference nor the shim.
"""
import io
import contextlib
import logging
import math
import os
import sys
import warnings

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT = os.environ.get("KGE_GOLDEN_OUT", HERE)        # tests/test_golden_oracle.py regenerates into a scratch directory
sys.path.insert(0, os.path.join(ROOT, "oracle", "tf_shim"))
sys.path.insert(0, "/root/reference")
sys.path.insert(0, ROOT)
warnings.filterwarnings("ignore", category=SyntaxWarning)

import numpy as np                      # noqa: E402
import tensorflow as tf                 # noqa: E402  (the shim)
from KGE import score as S, loss as Ls, constraint as Cn   # noqa: E402  (the reference)
from KGE.models.translating_based.UM import UM              # noqa: E402
from KGE.models.translating_based.SE import SE              # noqa: E402
from KGE.models.translating_based.TransE import TransE      # noqa: E402
from KGE.models.translating_based.TransH import TransH      # noqa: E402
from KGE.models.translating_based.TransR import TransR      # noqa: E402
from KGE.models.translating_based.TransD import TransD      # noqa: E402
from KGE.models.translating_based.RotatE import RotatE      # noqa: E402
from KGE.models.semantic_based.DistMult import DistMult     # noqa: E402
from KGE.models.semantic_based.RESCAL import RESCAL         # noqa: E402
from KGE.ns_strategy import TypedStrategy                   # noqa: E402
from oracle import philox                                   # noqa: E402  (inputs only: negatives are stored)

logging.getLogger().setLevel(logging.ERROR)
assert tf.__version__.endswith("shim")

N, R = 40, 5
MODELS = {"UM": UM, "SE": SE, "TransE": TransE, "TransH": TransH, "TransR": TransR, "TransD": TransD,
          "RotatE": RotatE, "DistMult": DistMult, "RESCAL": RESCAL}


def f32(a):
    return np.asarray(a, dtype=np.float32)


def npy(t):
    return np.asarray(t.numpy() if hasattr(t, "numpy") else t)


# ---------------------------------------------------------------------------------
# 1. plug-in ops: KGE/score.py, KGE/loss.py, KGE/constraint.py
# ---------------------------------------------------------------------------------
def gen_ops():
    rng = np.random.default_rng(20240501)
    out = {}
    x, y = f32(rng.uniform(-1, 1, (10, 5))), f32(rng.uniform(-1, 1, (10, 5)))
    xi, yi = f32(rng.uniform(-1, 1, (10, 5))), f32(rng.uniform(-1, 1, (10, 5)))
    y[3] = x[3]                                     # exercises the 1e-9 clip
    out.update(x=x, y=y, xi=xi, yi=yi)
    for name, fn in [("lp1", S.LpDistance(1)), ("lp2", S.LpDistance(2)), ("lp3", S.LpDistance(3)),
                     ("lpinf", S.LpDistance(np.inf)), ("lppow2", S.LpDistancePow(2)),
                     ("lppow1", S.LpDistancePow(1)), ("dot", S.Dot())]:
        out["score_" + name] = npy(fn(tf.constant(x), tf.constant(y)))
        if name != "dot":
            out["cscore_" + name] = npy(fn(tf.complex(x, xi), tf.complex(y, yi)))
    pos, neg = f32(rng.uniform(-2, 2, 10)), f32(rng.uniform(-2, 2, 40))
    out.update(pos=pos, neg=neg)
    for name, fn in [("hinge", Ls.PairwiseHingeLoss(margin=1.0)), ("hinge3", Ls.PairwiseHingeLoss(margin=3.0)),
                     ("logistic", Ls.PairwiseLogisticLoss()), ("bce", Ls.BinaryCrossEntropyLoss()),
                     ("selfadv", Ls.SelfAdversarialNegativeSamplingLoss(margin=3.0, temperature=1.0)),
                     ("selfadv2", Ls.SelfAdversarialNegativeSamplingLoss(margin=1.5, temperature=0.5)),
                     ("square", Ls.SquareErrorLoss())]:
        out["loss_" + name] = npy(fn(tf.constant(pos), tf.constant(neg)))
    X = f32(rng.uniform(0, 1, (10, 5)))
    X[2] *= 0.1
    out["X"] = X
    for p, tag in [(1, "1"), (2, "2"), (np.inf, "inf")]:
        out["norm_p" + tag] = npy(Cn.normalized_embeddings(tf.constant(X), p=p, value=1.0, axis=-1))
        out["clip_p" + tag] = npy(Cn.clip_constraint(tf.constant(X), p=p, value=1.0, axis=-1))
    out["soft_p2"] = npy(Cn.soft_constraint(tf.constant(X), p=2, value=1.0, axis=-1))
    out["lpreg_p2"] = npy(Cn.Lp_regularization(tf.constant(X), p=2, axis=-1))
    out["lpreg_p3"] = npy(Cn.Lp_regularization(tf.constant(X), p=3, axis=-1))
    np.savez_compressed(os.path.join(OUT, "ops.npz"), **out)
    return len(out)


# ---------------------------------------------------------------------------------
# 2. models: score_hrt, __run_single_batch, get_rank / evaluate
# ---------------------------------------------------------------------------------
def table_shapes(model, d, dr):
    return {
        "UM": {"ent_emb": (N, d)},
        "SE": {"ent_emb": (N, d), "rel_proj_h": (R, d, d), "rel_proj_t": (R, d, d)},
        "TransE": {"ent_emb": (N, d), "rel_emb": (R, d)},
        "TransH": {"ent_emb": (N, d), "rel_emb": (R, d), "rel_hyper": (R, d)},
        "TransR": {"ent_emb": (N, d), "rel_emb": (R, dr), "rel_proj": (R, d, dr)},
        "TransD": {"ent_emb": (N, d), "ent_proj": (N, d), "rel_emb": (R, dr), "rel_proj": (R, dr)},
        "RotatE": {"ent_emb": (N, d, 2), "rel_emb": (R, d)},
        "DistMult": {"ent_emb": (N, d), "rel_inter": (R, d)},
        "RESCAL": {"ent_emb": (N, d), "rel_inter": (R, d, d)},
    }[model]


def random_weights(model, d, dr, rng, rot_limit):
    out = {}
    for k, shp in table_shapes(model, d, dr).items():
        lim = rot_limit if model == "RotatE" else math.sqrt(6.0 / shp[-1])
        out[k] = f32(rng.uniform(-lim, lim, shp))
    return out


class GoldenSampler:
    """Sampler *instance* (passed through untouched, BaseModel.py:256-258): serves the
    Philox-spec negatives for (seed, step) so the fixture is reproducible by the kernels."""

    def __init__(self, seed, pools=None):
        self.seed, self.step, self.pools = seed, 0, pools

    def __call__(self, X, negative_ratio, side):
        Xn = npy(X)
        if self.pools is not None:
            ids = philox.sample_typed(self.seed, self.step, side, Xn, negative_ratio, self.pools)
        else:
            ids = philox.sample_uniform(self.seed, self.step, side, len(Xn), negative_ratio, N)
        return tf.constant(ids, dtype=X.dtype)


def build_reference_model(model, score, loss, k, side, constraint, cw, d, dr, w0, sampler, optimizer=None):
    emb = {"embedding_size": d} if model not in ("TransR", "TransD") else \
        {"ent_embedding_size": d, "rel_embedding_size": dr}
    kw = dict(embedding_params=emb, negative_ratio=k, corrupt_side=side, loss_fn=loss, ns_strategy=sampler)
    if model not in ("DistMult", "RESCAL"):
        kw["score_fn"] = score
    if model != "RotatE":
        kw["constraint"] = constraint
    if model in ("TransH", "DistMult", "RESCAL"):
        kw["constraint_weight"] = cw
    m = MODELS[model](**kw)
    m.metadata = {"ind2ent": list(range(N)), "ind2rel": list(range(R))}
    W = {kk: tf.Variable(v) for kk, v in w0.items()}
    if model == "RotatE" or (model == "TransE" and constraint):
        # model_weights_initial is unusable for these two: set the dict directly
        m.model_weights = W
        if model == "RotatE":
            m.limit = (getattr(loss, "margin", 6.0) + 2.0) / d
    else:
        m._model_weights_initial = W
        m._init_embeddings(seed=None)        # the reference's own path (RESCAL normalises here)
    # BaseModel.py:243-246: "Adam" -> tf.optimizers.Adam(), anything else is used as given
    m._KGEModel__optimizer = tf.optimizers.Adam() if optimizer is None else optimizer
    return m


SCORES = {"l1": lambda: S.LpDistance(1), "l2": lambda: S.LpDistance(2), "linf": lambda: S.LpDistance(np.inf),
          "l2pow": lambda: S.LpDistancePow(2), "dot": lambda: S.Dot(), None: lambda: None}
LOSSES = {"hinge": lambda: Ls.PairwiseHingeLoss(margin=1.0), "logistic": lambda: Ls.PairwiseLogisticLoss(),
          "bce": lambda: Ls.BinaryCrossEntropyLoss(),
          "selfadv": lambda: Ls.SelfAdversarialNegativeSamplingLoss(margin=3.0, temperature=1.0),
          "square": lambda: Ls.SquareErrorLoss()}

# name: (model, score, loss, k, side, constraint, cw, d, d_rel)
CASES = {
    "um": ("UM", "l2pow", "hinge", 4, "h+t", True, 1.0, 8, 8),
    "se": ("SE", "l1", "hinge", 2, "t", True, 1.0, 6, 6),
    "transe": ("TransE", "l2", "hinge", 4, "h+t", True, 1.0, 8, 8),
    "transe_l1_logistic_h": ("TransE", "l1", "logistic", 3, "h", False, 1.0, 8, 8),
    "transe_dot_square": ("TransE", "dot", "square", 2, "t", True, 1.0, 8, 8),
    "transh": ("TransH", "l2pow", "hinge", 4, "h+t", True, 0.5, 8, 8),
    "transh_nocon": ("TransH", "l2pow", "hinge", 4, "h+t", False, 1.0, 8, 8),
    "transr": ("TransR", "l2pow", "hinge", 4, "h+t", True, 1.0, 8, 6),
    "transr_nocon": ("TransR", "l2", "bce", 2, "h", False, 1.0, 6, 8),
    "transd": ("TransD", "l2pow", "hinge", 4, "h+t", True, 1.0, 8, 6),
    "transd_wide_rel": ("TransD", "l2pow", "hinge", 4, "h+t", True, 1.0, 6, 8),
    "rotate": ("RotatE", "l1", "selfadv", 6, "h+t", True, 1.0, 7, 7),
    "rotate_l2_bce": ("RotatE", "l2", "bce", 2, "t", True, 1.0, 7, 7),
    "distmult": ("DistMult", None, "bce", 6, "h+t", True, 1.0, 8, 8),
    "distmult_hinge_nocon": ("DistMult", None, "hinge", 2, "h", False, 1.0, 8, 8),
    "rescal": ("RESCAL", None, "square", 4, "h+t", True, 1.0, 6, 6),
    "rescal_logistic_nocon": ("RESCAL", None, "logistic", 2, "t", False, 1.0, 6, 6),
    # round 2: one more per model, other sizes (rectangular TransR / TransD both ways), sides and score / loss pairings
    "extra_transe": ("TransE", "l1", "logistic", 3, "t", True, 1.0, 12, 12),
    "extra_distmult": ("DistMult", None, "hinge", 5, "h+t", True, 0.5, 10, 10),
    "extra_transr": ("TransR", "l2", "hinge", 2, "h", True, 1.0, 6, 9),
    "extra_rotate": ("RotatE", "l2", "selfadv", 4, "h+t", True, 1.0, 10, 10),
    "extra_transh": ("TransH", "l2pow", "bce", 3, "h+t", True, 0.25, 8, 8),
    "extra_transd": ("TransD", "l1", "square", 2, "t", False, 1.0, 7, 11),
    "extra_rescal": ("RESCAL", None, "logistic", 2, "h+t", True, 1.0, 6, 6),
    "extra_se": ("SE", "l2", "hinge", 3, "h+t", True, 1.0, 5, 5),
    "extra_um": ("UM", "l1", "hinge", 2, "h", True, 1.0, 9, 9),
}
# user-supplied Keras optimizers (BaseModel.py:243-246).  name: (case spec, optimizer, lr)
OPT_CASES = {
    "opt_transe_sgd": (("TransE", "l1", "hinge", 4, "h+t", True, 1.0, 8, 8), "sgd", 0.05),
    "opt_transe_adagrad": (("TransE", "l1", "hinge", 4, "h+t", True, 1.0, 8, 8), "adagrad", 0.05),
    "opt_distmult_sgd": (("DistMult", None, "bce", 6, "h+t", True, 1.0, 8, 8), "sgd", 0.05),
    "opt_distmult_adagrad": (("DistMult", None, "bce", 6, "h+t", True, 1.0, 8, 8), "adagrad", 0.05),
    "opt_transh_sgd": (("TransH", "l2pow", "hinge", 4, "h+t", True, 0.5, 8, 8), "sgd", 0.05),       # dense regulariser
    "opt_transh_adagrad": (("TransH", "l2pow", "hinge", 4, "h+t", True, 0.5, 8, 8), "adagrad", 0.05),
    "opt_transr_sgd": (("TransR", "l2pow", "hinge", 4, "h+t", True, 1.0, 8, 6), "sgd", 0.05),       # matrix table
    "opt_transr_adagrad": (("TransR", "l2pow", "hinge", 4, "h+t", True, 1.0, 8, 6), "adagrad", 0.05),
    "opt_rotate_adagrad": (("RotatE", "l1", "selfadv", 6, "h+t", True, 1.0, 7, 7), "adagrad", 0.05),
    "opt_transd_sgd": (("TransD", "l2pow", "hinge", 4, "h+t", True, 1.0, 8, 6), "sgd", 0.05),
}
B, STEPS, SEED = 16, 3, 4242


def make_optimizer(kind, lr):
    if kind == "sgd":
        return tf.optimizers.SGD(learning_rate=lr)
    if kind == "adagrad":
        return tf.optimizers.Adagrad(learning_rate=lr, initial_accumulator_value=0.1, epsilon=1e-7)
    raise ValueError(kind)


def gen_case(name, spec, opt=None):
    model, sc, ls, k, side, con, cw, d, dr = spec
    rng = np.random.default_rng(abs(hash(name)) % (2 ** 31) if False else sum(map(ord, name)))
    loss = LOSSES[ls]()
    rot_limit = (getattr(loss, "margin", 6.0) + 2.0) / d
    w_in = random_weights(model, d, dr, rng, rot_limit)
    sampler = GoldenSampler(SEED)
    m = build_reference_model(model, SCORES[sc](), loss, k, side, con, cw, d, dr, w_in, sampler,
                              optimizer=None if opt is None else make_optimizer(*opt))
    out = {"w0_" + kk: npy(v).copy() for kk, v in m.model_weights.items()}      # post-init state

    def triples(n):
        return np.stack([rng.integers(0, N, n), rng.integers(0, R, n), rng.integers(0, N, n)], 1).astype(np.int64)

    # -- score_hrt on id vectors and on the h=None / t=None expansions (BaseModel.py:410-430)
    Xs = triples(12)
    out["score_X"] = Xs
    out["score_batch"] = npy(m.score_hrt(tf.constant(Xs[:, 0]), tf.constant(Xs[:, 1]), tf.constant(Xs[:, 2])))
    q = tf.constant(Xs[0])
    out["score_all_h"] = npy(m.score_hrt(h=None, r=q[1], t=q[2]))
    out["score_all_t"] = npy(m.score_hrt(h=q[0], r=q[1], t=None))

    # -- steps: train, train, validation, train  (BaseModel.py:137-145 interleaving)
    rec = {"X": [], "neg": [], "loss": [], "is_train": []}
    orig_ns = m._KGEModel__negative_sampling

    def recording_ns(X, strategy):
        neg = orig_ns(X, strategy)
        rec["neg"].append(npy(neg).copy())
        return neg
    m._KGEModel__negative_sampling = recording_ns
    plan = [True, True, False, True]
    for step, is_train in enumerate(plan):
        X = triples(B)
        sampler.step = step
        l = m._KGEModel__run_single_batch(batch_data=tf.constant(X), is_train=is_train)
        rec["X"].append(X); rec["loss"].append(float(l)); rec["is_train"].append(is_train)
        if step == 0:
            out.update({"w1_" + kk: npy(v).copy() for kk, v in m.model_weights.items()})
    out["step_X"] = np.stack(rec["X"])
    out["step_neg"] = np.stack(rec["neg"])                 # (steps, B*k_eff, 3) reference row order
    out["step_loss"] = np.asarray(rec["loss"], dtype=np.float64)
    out["step_is_train"] = np.asarray(rec["is_train"])
    out.update({"wT_" + kk: npy(v).copy() for kk, v in m.model_weights.items()})

    # -- ranks on the trained weights (BaseModel.py:620-654) and evaluate() (:578-618)
    pos = triples(300)
    ev = pos[:10].copy()
    ev[::4] = triples(len(ev[::4]))
    out["rank_pos"], out["rank_X"] = pos, ev
    for s in ("h", "t"):
        out["rank_raw_" + s] = np.asarray([m.get_rank(tf.constant(x[None]), None, s) for x in ev], dtype=np.int64)
        out["rank_flt_" + s] = np.asarray([m.get_rank(tf.constant(x[None]), pos, s) for x in ev], dtype=np.int64)
        # the two score paths get_rank compares (BaseModel.py:639-654): every candidate through the
        # h=None / t=None expansion, the true triple through the scalar call — the tests take the
        # documented tie band (and the reference's occasional self-count) from THESE
        sc_all = []
        for x in ev:
            q = tf.constant(x)
            sc_all.append(npy(m.score_hrt(h=None, r=q[1], t=q[2]) if s == "h" else m.score_hrt(h=q[0], r=q[1], t=None)))
        out["rank_scores_" + s] = np.stack(sc_all).astype(np.float32)
    out["rank_pos_score"] = np.asarray([float(npy(m.score_hrt(tf.constant(x)[0], tf.constant(x)[1], tf.constant(x)[2])).reshape(-1)[0])
                                        for x in ev], dtype=np.float32)
    with contextlib.redirect_stderr(io.StringIO()):
        res = m.evaluate(ev, "t", pos)
    out["eval_keys"] = np.asarray(list(res.keys()))
    out["eval_vals"] = np.asarray([float(v) for v in res.values()], dtype=np.float64)
    out["spec"] = np.asarray([str(x) for x in spec])
    if opt is not None:
        out["opt"] = np.asarray([str(opt[0]), repr(float(opt[1]))])
    np.savez_compressed(os.path.join(OUT, "case_%s.npz" % name), **out)
    return rec["loss"]


# ---------------------------------------------------------------------------------
# 3. TypedStrategy on the reference's own (NumPy global RNG) stream: invariants only
# ---------------------------------------------------------------------------------
def gen_typed():
    rng = np.random.default_rng(99)
    ind2type = [["A", "B", "C"][i] for i in rng.integers(0, 3, N)]
    type2inds = {t: np.array([i for i, ti in enumerate(ind2type) if ti == t]) for t in np.unique(ind2type)}
    X = np.stack([rng.integers(0, N, 64), rng.integers(0, R, 64), rng.integers(0, N, 64)], 1).astype(np.int64)
    np.random.seed(7)
    strat = TypedStrategy(pool=None, metadata={"type2inds": type2inds, "ind2type": ind2type})
    out = {"ind2type": np.asarray(ind2type), "X": X}
    for side in ("h", "t"):
        out["neg_" + side] = npy(strat(tf.constant(X), negative_ratio=5, side=side))
    np.savez_compressed(os.path.join(OUT, "typed.npz"), **out)


# ---------------------------------------------------------------------------------
# 4. the reference's full train() + evaluate() loop (BaseModel.py:58-190, 578-618)
# ---------------------------------------------------------------------------------
def gen_train_loop():
    rng = np.random.default_rng(31337)
    d, k, bs, epochs = 8, 4, 8, 2
    train = np.stack([rng.integers(0, N, 50), rng.integers(0, R, 50), rng.integers(0, N, 50)], 1).astype(np.int64)
    val = np.stack([rng.integers(0, N, 12), rng.integers(0, R, 12), rng.integers(0, N, 12)], 1).astype(np.int64)
    w_in = random_weights("TransH", d, d, rng, None)

    class Recording(GoldenSampler):
        """one Philox step per __run_single_batch; logs the batch stream train() produced"""
        def __init__(self):
            super().__init__(SEED)
            self.batches, self.calls = [], 0

        def __call__(self, X, negative_ratio, side):
            if self.calls % 2 == 0:          # 'h+t': h call then t call per batch
                self.step = self.calls // 2
                self.batches.append(npy(X).copy())
            self.calls += 1
            return super().__call__(X, negative_ratio, side)

    smp = Recording()
    m = TransH({"embedding_size": d}, k, "h+t", S.LpDistancePow(2), Ls.PairwiseHingeLoss(margin=1.0), smp,
               constraint=False, constraint_weight=1.0)
    losses = []
    orig = m._KGEModel__run_single_batch

    def rec_run(batch_data, is_train):
        l = orig(batch_data=batch_data, is_train=is_train)
        losses.append((float(l), bool(is_train)))
        return l
    m._KGEModel__run_single_batch = rec_run
    md = {"ind2ent": list(range(N)), "ind2rel": list(range(R))}
    with contextlib.redirect_stderr(io.StringIO()):
        m.train(train, val, md, epochs=epochs, batch_size=bs, seed=5,
                model_weights_initial={kk: tf.Variable(v) for kk, v in w_in.items()},
                log_path="/tmp/kge_golden_logs")
        res = m.evaluate(val, "h", np.concatenate([train, val]))
    out = {"w0_" + kk: v for kk, v in w_in.items()}
    out.update({"wT_" + kk: npy(v).copy() for kk, v in m.model_weights.items()})
    out["batches"] = np.stack(smp.batches)
    out["loss"] = np.asarray([l for l, _ in losses]); out["is_train"] = np.asarray([t for _, t in losses])
    out["train"], out["val"] = train, val
    out["eval_keys"] = np.asarray(list(res.keys()))
    out["eval_vals"] = np.asarray([float(v) for v in res.values()])
    out["cfg"] = np.asarray([d, k, bs, epochs, SEED])
    np.savez_compressed(os.path.join(OUT, "train_loop_transh.npz"), **out)
    return len(smp.batches)


if __name__ == "__main__":
    print("ops:", gen_ops(), "arrays")
    for name, spec in CASES.items():
        ls = gen_case(name, spec)
        print("case %-24s losses %s" % (name, " ".join("%.5f" % v for v in ls)))
    for name, (spec, kind, lr) in OPT_CASES.items():
        ls = gen_case(name, spec, (kind, lr))
        print("case %-24s losses %s" % (name, " ".join("%.5f" % v for v in ls)))
    gen_typed()
    print("train loop batches:", gen_train_loop())
