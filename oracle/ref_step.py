"""Training step, ranking and metrics — torch-CPU restatement (TEST INFRASTRUCTURE).

Follows KGE/models/base_model/BaseModel.py:293-408 (step), :620-654 (rank),
:597-616 (evaluate), KGE/metrics.py:5-25, and the TensorFlow 2.5 semantics of
SURVEY.md Appendix B (IndexedSlices concatenation, clip_by_norm over un-merged
values, Keras Adam sparse = dedup-sum + DENSE decay/update).  The backward is
autograd, independent of the hand-derived CUDA gradients.
"""
import math
import numpy as np
import torch
from . import philox
from .ref_models import Weights


class TFAdam:
    """tf.optimizers.Adam() defaults (App. B.2); state keyed by table name."""

    def __init__(self, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-7):
        self.lr, self.b1, self.b2, self.eps = lr, beta1, beta2, eps
        self.t = 0
        self.m, self.v = {}, {}

    def _slots(self, k, w):
        if k not in self.m:
            self.m[k] = torch.zeros_like(w)
            self.v[k] = torch.zeros_like(w)
        return self.m[k], self.v[k]

    def begin_step(self):
        self.t += 1
        f32 = np.float32
        # Keras computes the coefficients in the variable dtype (fp32)
        b1p, b2p = f32(self.b1) ** f32(self.t), f32(self.b2) ** f32(self.t)
        self.lr_t = float(f32(self.lr) * np.sqrt(f32(1) - b2p) / (f32(1) - b1p))

    def apply_dense(self, k, w, g):
        m, v = self._slots(k, w)
        m.mul_(self.b1).add_(g, alpha=1 - self.b1)
        v.mul_(self.b2).add_(g * g, alpha=1 - self.b2)
        w.sub_(self.lr_t * m / (torch.sqrt(v) + self.eps))

    def apply_sparse(self, k, w, idx, vals):
        m, v = self._slots(k, w)
        uniq, inv = torch.unique(idx, return_inverse=True)
        summed = torch.zeros((len(uniq),) + tuple(vals.shape[1:]), dtype=vals.dtype)
        summed.index_add_(0, inv, vals)            # sequential in occurrence order
        m.mul_(self.b1)
        m[uniq] += (1 - self.b1) * summed
        v.mul_(self.b2)
        v[uniq] += (1 - self.b2) * summed * summed
        w.sub_(self.lr_t * m / (torch.sqrt(v) + self.eps))


class TFSGD:
    """tf.optimizers.SGD(momentum=0), TF 2.5: dense var -= lr g; sparse = scatter_add(var, idx, -lr g) per
    occurrence (gradient_descent.py _resource_apply_sparse_duplicate_indices)."""

    def __init__(self, lr=0.01):
        self.lr, self.t = lr, 0

    def begin_step(self):
        self.t += 1

    def apply_dense(self, k, w, g):
        w.sub_(self.lr * g)

    def apply_sparse(self, k, w, idx, vals):
        w.index_add_(0, idx, -self.lr * vals)


class TFAdagrad:
    """tf.optimizers.Adagrad, TF 2.5: accumulator starts at initial_accumulator_value; sparse apply on the
    de-duplicated (summed) rows: acc[i] += g^2; var[i] -= lr g / (sqrt(acc[i]) + eps)
    (optimizer_v2._resource_apply_sparse_duplicate_indices + ResourceSparseApplyAdagradV2)."""

    def __init__(self, lr=0.001, initial_accumulator_value=0.1, eps=1e-7):
        self.lr, self.init, self.eps, self.t, self.acc = lr, initial_accumulator_value, eps, 0, {}

    def begin_step(self):
        self.t += 1

    def _acc(self, k, w):
        if k not in self.acc:
            self.acc[k] = torch.full_like(w, self.init)
        return self.acc[k]

    def apply_dense(self, k, w, g):
        a = self._acc(k, w)
        a.add_(g * g)
        w.sub_(self.lr * g / (torch.sqrt(a) + self.eps))

    def apply_sparse(self, k, w, idx, vals):
        a = self._acc(k, w)
        uniq, inv = torch.unique(idx, return_inverse=True)
        summed = torch.zeros((len(uniq),) + tuple(vals.shape[1:]), dtype=vals.dtype)
        summed.index_add_(0, inv, vals)
        a[uniq] += summed * summed
        w[uniq] -= self.lr * summed / (torch.sqrt(a[uniq]) + self.eps)


def clip_by_norm(values, clip):
    """tf.clip_by_norm (BaseModel.py:327): g*clip/max(||g||, clip)."""
    l2 = torch.sqrt(torch.sum(values * values))
    return values * clip / torch.maximum(l2, torch.tensor(clip, dtype=values.dtype))


def build_negatives(X, neg_ids, k_eff, corrupt_side):
    """(B*k_eff, 3) negative triples in reference row order (BaseModel.py:360-408,354-356)."""
    X = torch.as_tensor(X, dtype=torch.int64)
    ids = torch.as_tensor(np.asarray(neg_ids), dtype=torch.int64)
    rep = torch.repeat_interleave(X, k_eff, dim=0)
    h, r, t = rep[:, 0].clone(), rep[:, 1], rep[:, 2].clone()
    c = torch.arange(len(ids)) % k_eff
    if corrupt_side == "h":
        is_h = torch.ones_like(c, dtype=torch.bool)
    elif corrupt_side == "t":
        is_h = torch.zeros_like(c, dtype=torch.bool)
    else:
        is_h = (c % 2 == 0)
    h[is_h] = ids[is_h]
    t[~is_h] = ids[~is_h]
    return torch.stack([h, r, t], dim=1)


def run_single_batch(model, W, opt, X, neg_ids, k_eff, corrupt_side, is_train, clip=5.0):
    """One __run_single_batch (BaseModel.py:293-330). Mutates W (and opt). Returns the
    loss as a python float plus a dict of diagnostics used by parity tests."""
    X = torch.as_tensor(np.asarray(X), dtype=torch.int64)
    neg = build_negatives(X, neg_ids, k_eff, corrupt_side)
    W.start_tape()
    c = model.constraint_loss(W, X)
    pos_s = model.score_hrt(W, X[:, 0], X[:, 1], X[:, 2])
    neg_s = model.score_hrt(W, neg[:, 0], neg[:, 1], neg[:, 2])
    loss = model.loss_fn(pos_s, neg_s) + c
    tape = W.stop_tape()
    diag = {"pos_score": pos_s.detach().clone(), "neg_score": neg_s.detach().clone(), "norm2": {}}
    if is_train:
        leaves = [rows for (_, _, rows) in tape["sparse"]] + list(tape["dense"].values())
        grads = torch.autograd.grad(loss, leaves, allow_unused=True)
        ns = len(tape["sparse"])
        per_var = {}
        for (k, idx, rows), g in zip(tape["sparse"], grads[:ns]):
            if g is None:
                g = torch.zeros_like(rows)
            per_var.setdefault(k, {"idx": [], "val": [], "dense": None})
            per_var[k]["idx"].append(idx.reshape(-1))
            per_var[k]["val"].append(g.reshape((-1,) + tuple(W.data[k].shape[1:])))
        for k, g in zip(tape["dense"].keys(), grads[ns:]):
            per_var.setdefault(k, {"idx": [], "val": [], "dense": None})
            per_var[k]["dense"] = g if g is not None else torch.zeros_like(W.data[k])
        opt.begin_step()
        for k in W.data:                      # variable order = model_weights order
            if k not in per_var:
                continue
            pv = per_var[k]
            if pv["dense"] is not None:
                g = pv["dense"].clone()
                for idx, val in zip(pv["idx"], pv["val"]):
                    g.index_add_(0, idx, val)
                diag["norm2"][k] = float(torch.sum(g * g))
                opt.apply_dense(k, W.data[k], clip_by_norm(g, clip))
            else:
                idx = torch.cat(pv["idx"])
                val = torch.cat(pv["val"])
                diag["norm2"][k] = float(torch.sum(val * val))
                opt.apply_sparse(k, W.data[k], idx, clip_by_norm(val, clip))
    return float(loss.detach()), diag


def get_scores_all(model, W, x, side):
    """score_hrt(h=None,...) / (...,t=None) expansion (BaseModel.py:420-430, 639-644)."""
    n = model.n_ent
    ar = torch.arange(n, dtype=torch.int64)
    h, r, t = (torch.full((n,), int(v), dtype=torch.int64) for v in x)
    with torch.no_grad():
        if side == "h":
            return model.score_hrt(W, ar, r, t)
        return model.score_hrt(W, h, r, ar)


def get_rank(model, W, x, positive_X, side):
    """BaseModel.py:620-654; accumulator widened from int16 to int64."""
    x = np.asarray(x).reshape(3)
    scores = get_scores_all(model, W, x, side).clone()
    fs, cs = (2, 0) if side == "h" else (0, 2)
    if positive_X is not None:
        P = np.asarray(positive_X)
        sel = P[(P[:, 1] == x[1]) & (P[:, fs] == x[fs]), cs]
        scores[torch.as_tensor(sel, dtype=torch.int64)] = -math.inf
    with torch.no_grad():
        one = lambda v: torch.tensor([int(v)], dtype=torch.int64)
        pos = model.score_hrt(W, one(x[0]), one(x[1]), one(x[2]))[0]
    return int(torch.sum(scores > pos)) + 1


def metrics(ranks):
    """KGE/metrics.py:5-25 and the dict of BaseModel.py:606-616."""
    r = np.asarray(ranks, dtype=np.float64)
    return {
        "mean_rank": float(np.mean(r)),
        "mean_reciprocal_rank": float(np.mean(1 / r)),
        "median_rank": float(np.median(r)),
        "geometric_mean_rank": float(np.exp(np.mean(np.log(r)))),
        "harmonic_mean_rank": float(len(r) / np.sum(1 / r)),
        "std_rank": float(np.std(r)),
        "hit@1": float(np.mean(r <= 1)),
        "hit@3": float(np.mean(r <= 3)),
        "hit@10": float(np.mean(r <= 10)),
    }


def evaluate(model, W, eval_X, side, positive_X=None):
    ranks = [get_rank(model, W, x, positive_X, side) for x in np.asarray(eval_X)]
    return metrics(ranks), ranks


def sample_negatives(seed, step, X, negative_ratio, corrupt_side, n_ent, pool=None, typed=None):
    return philox.negatives(seed, step, X, negative_ratio, corrupt_side, n_ent, pool, typed)
