"""Scores, losses, constraints — torch-CPU fp32 restatement (TEST INFRASTRUCTURE).

Each function follows the reference line it cites, op for op, so that autograd
reproduces the TensorFlow gradient conventions of SURVEY.md Appendix B.4:
``clip_by_value`` passes gradient on the closed interval, ``abs`` has gradient
``sign`` (0 at 0), complex ``abs`` has gradient ``z/|z|`` (0 at 0).
"""
import math
import torch


class _ClipByValue(torch.autograd.Function):
    """tf.clip_by_value: gradient passes where lo <= x <= hi (inclusive)."""

    @staticmethod
    def forward(ctx, x, lo, hi):
        ctx.save_for_backward((x >= lo) & (x <= hi))
        return x.clamp(lo, hi)

    @staticmethod
    def backward(ctx, g):
        (mask,) = ctx.saved_tensors
        return g * mask.to(g.dtype), None, None


def clip_by_value(x, lo, hi):
    return _ClipByValue.apply(x, lo, hi)


class _ComplexAbs(torch.autograd.Function):
    """tf.abs on (re, im): modulus; gradient z/|z|, 0 at z == 0."""

    @staticmethod
    def forward(ctx, re, im):
        mod = torch.sqrt(re * re + im * im)
        ctx.save_for_backward(re, im, mod)
        return mod

    @staticmethod
    def backward(ctx, g):
        re, im, mod = ctx.saved_tensors
        safe = torch.where(mod > 0, mod, torch.ones_like(mod))
        nz = (mod > 0).to(g.dtype)
        return g * re / safe * nz, g * im / safe * nz


def _abs_diff(x, y):
    """|x - y| elementwise; x, y real tensors or (re, im) tuples (RotatE)."""
    if isinstance(x, tuple):
        return _ComplexAbs.apply(x[0] - y[0], x[1] - y[1])
    return torch.abs(x - y)


# ---- scores: KGE/score.py -------------------------------------------------------
class LpDistance:
    """score.py:49-63."""

    def __init__(self, p):
        self.p = p

    def __call__(self, x, y):
        a = _abs_diff(x, y)
        if self.p == math.inf:
            # tf.reduce_max gradient splits equally among ties -> torch.amax does too
            return -torch.amax(a, dim=-1)
        s = torch.sum(torch.pow(a, self.p), dim=-1)
        return -torch.pow(clip_by_value(s, 1e-9, math.inf), 1.0 / self.p)


class LpDistancePow:
    """score.py:65-76."""

    def __init__(self, p):
        self.p = p

    def __call__(self, x, y):
        return -torch.pow(LpDistance(self.p)(x, y), 2)


class Dot:
    """score.py:78-89."""

    def __call__(self, x, y):
        return torch.sum(x * y, dim=-1)


# ---- losses: KGE/loss.py --------------------------------------------------------
class PairwiseHingeLoss:
    """loss.py:76-82 (divides by the *repeated* length n*k)."""

    def __init__(self, margin):
        self.margin = margin

    def __call__(self, pos, neg):
        k = neg.shape[0] // pos.shape[0]
        pos = torch.repeat_interleave(pos, k)
        return torch.sum(clip_by_value(self.margin + neg - pos, 0.0, math.inf)) / pos.shape[0]


class PairwiseLogisticLoss:
    """loss.py:107-113 (naive log(1+exp), no normalisation)."""

    def __call__(self, pos, neg):
        k = neg.shape[0] // pos.shape[0]
        pos = torch.repeat_interleave(pos, k)
        return torch.sum(torch.log(1 + torch.exp(neg - pos)))


class BinaryCrossEntropyLoss:
    """loss.py:135-143."""

    def __call__(self, pos, neg):
        F = torch.nn.functional
        return -(torch.sum(F.logsigmoid(pos)) + torch.sum(F.logsigmoid(-neg))) / pos.shape[0]


class SelfAdversarialNegativeSamplingLoss:
    """loss.py:170-182."""

    def __init__(self, margin, temperature):
        self.margin = margin
        self.temperature = temperature

    def __call__(self, pos, neg):
        F = torch.nn.functional
        neg = neg.reshape(pos.shape[0], neg.shape[0] // pos.shape[0])
        prob = torch.softmax(self.temperature * neg, dim=-1).detach()
        pos_ll = torch.sum(F.logsigmoid(pos + self.margin))
        neg_ll = torch.sum(prob * F.logsigmoid(-neg - self.margin))
        return -(pos_ll + neg_ll) / pos.shape[0]


class SquareErrorLoss:
    """loss.py:198-204."""

    def __call__(self, pos, neg):
        return (torch.sum((pos - 1.0) ** 2) + torch.sum(neg ** 2)) / 2 / pos.shape[0]


# ---- constraints: KGE/constraint.py ---------------------------------------------
def _norm(X, p, axis):
    """constraint.py:26-29 (same expression in 62-65 and 93-96)."""
    if p == math.inf:
        return torch.amax(torch.abs(X), dim=axis, keepdim=True)
    return torch.pow(torch.sum(torch.pow(torch.abs(X), p), dim=axis, keepdim=True), 1 / p)


def normalized_embeddings(X, p, value, axis):
    """constraint.py:4-31 (no epsilon: zero row -> NaN)."""
    return X / _norm(X, p, axis) * value


def soft_constraint(X, p, value, axis):
    """constraint.py:34-67."""
    return torch.sum(clip_by_value(torch.pow(_norm(X, p, axis), p) - value, 0.0, math.inf))


def clip_constraint(X, p, value, axis):
    """constraint.py:70-99 (mask is a cast of a comparison: no gradient through it)."""
    norm = _norm(X, p, axis)
    mask = (norm < value).to(X.dtype)
    return mask * X + (1 - mask) * (X / clip_by_value(norm, 1e-9, math.inf) * value)


def Lp_regularization(X, p, axis):
    """constraint.py:102-126."""
    return torch.sum(torch.pow(torch.abs(X), p), dim=axis)
