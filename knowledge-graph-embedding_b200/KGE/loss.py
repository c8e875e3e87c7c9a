"""Loss descriptors — drop-in for KGE/loss.py of the reference (PairwiseHingeLoss :49,
PairwiseLogisticLoss :84, BinaryCrossEntropyLoss :115,
SelfAdversarialNegativeSamplingLoss :145, SquareErrorLoss :184).

During training the class and its `margin` / `temperature` select the loss branch of
the fused step kernel (csrc/loss.cuh); `__call__(pos_score, neg_score)` is the
standalone form (tests/test_loss.py) written with torch ops.  The three different
normalisations of the reference (/(n*k), none, /n) are kept on purpose.
"""
import torch
import torch.nn.functional as F

from .score import _as_tensor


class Loss:
    def __init__(self):
        raise NotImplementedError("a Loss subclass must define its parameters in __init__()")

    def __call__(self, pos_score, neg_score):
        raise NotImplementedError("a Loss subclass must implement __call__(pos_score, neg_score)")


def _paired(pos_score, neg_score):
    pos, neg = _as_tensor(pos_score), _as_tensor(neg_score)
    per_pos = neg.shape[0] // pos.shape[0]
    return pos, neg, per_pos


class PairwiseHingeLoss(Loss):
    def __init__(self, margin):
        self.margin = margin

    def __call__(self, pos_score, neg_score):
        pos, neg, k = _paired(pos_score, neg_score)
        spread = pos.repeat_interleave(k)
        return (self.margin + neg - spread).clamp_min(0).sum() / spread.shape[0]


class PairwiseLogisticLoss(Loss):
    def __init__(self):
        pass

    def __call__(self, pos_score, neg_score):
        pos, neg, k = _paired(pos_score, neg_score)
        return torch.log1p(torch.exp(neg - pos.repeat_interleave(k))).sum()


class BinaryCrossEntropyLoss(Loss):
    def __init__(self):
        pass

    def __call__(self, pos_score, neg_score):
        pos, neg, _ = _paired(pos_score, neg_score)
        return -(F.logsigmoid(pos).sum() + F.logsigmoid(-neg).sum()) / pos.shape[0]


class SelfAdversarialNegativeSamplingLoss(Loss):
    def __init__(self, margin, temperature):
        self.margin = margin
        self.temperature = temperature

    def __call__(self, pos_score, neg_score):
        pos, neg, k = _paired(pos_score, neg_score)
        grid = neg.reshape(pos.shape[0], k)
        weight = torch.softmax(self.temperature * grid, dim=-1).detach()
        total = F.logsigmoid(pos + self.margin).sum() + (weight * F.logsigmoid(-grid - self.margin)).sum()
        return -total / pos.shape[0]


class SquareErrorLoss(Loss):
    def __init__(self):
        pass

    def __call__(self, pos_score, neg_score):
        pos, neg, _ = _paired(pos_score, neg_score)
        return ((pos - 1).pow(2).sum() + neg.pow(2).sum()) / 2 / pos.shape[0]
