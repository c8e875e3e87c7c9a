"""Score descriptors — drop-in for KGE/score.py of the reference (LpDistance :49,
LpDistancePow :65, Dot :78).

Inside `train()` / `evaluate()` / `score_hrt()` these objects are *descriptors*: their
class and `p` select the fused CUDA kernel variant (KGE/_b200/spec.py), nothing here
is called per batch.  `__call__` is kept for standalone use (the reference's
tests/test_score.py calls it on raw arrays): it evaluates the same formula with torch
elementwise ops on whatever device the inputs live on.
"""
import math

import numpy as np
import torch


def _as_tensor(x):
    if torch.is_tensor(x):
        return x
    a = np.asarray(x)
    if a.dtype == np.float64:
        a = a.astype(np.float32)
    return torch.as_tensor(a)


class Score:
    """Base class: subclasses define the parameters and `__call__(x, y) -> (n,)`."""

    def __init__(self):
        raise NotImplementedError("a Score subclass must define its parameters in __init__()")

    def __call__(self, x, y):
        raise NotImplementedError("a Score subclass must implement __call__(x, y)")


class LpDistance(Score):
    """-||x - y||_p ; p = inf gives -max|x - y| ; complex inputs use the modulus."""

    def __init__(self, p):
        self.p = p

    def __call__(self, x, y):
        gap = torch.abs(_as_tensor(x) - _as_tensor(y))
        if self.p == math.inf:
            return -gap.amax(dim=-1)
        total = gap.pow(self.p).sum(dim=-1).clamp_min(1e-9)
        return -total.pow(1.0 / self.p)


class LpDistancePow(Score):
    """-||x - y||_p ** 2."""

    def __init__(self, p):
        self.p = p

    def __call__(self, x, y):
        return -LpDistance(self.p)(x, y).pow(2)


SquaredLpDistance = LpDistancePow      # name used by BASELINE.json's prose


class Dot(Score):
    """<x, y>."""

    def __init__(self):
        pass

    def __call__(self, x, y):
        return (_as_tensor(x) * _as_tensor(y)).sum(dim=-1)
