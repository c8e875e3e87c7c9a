"""Constraint helpers — drop-in for KGE/constraint.py of the reference
(normalized_embeddings :4, soft_constraint :34, clip_constraint :70, Lp_regularization :102).

The per-batch constraints of the models run inside the CUDA library (k_constrain and
the dense-regulariser passes of k_update); these functions are the standalone forms
(tests/test_constraint.py), torch ops on the input's device.
"""
import math

import torch

from .score import _as_tensor


def _lp_norm(X, p, axis):
    mag = torch.abs(_as_tensor(X))
    if p == math.inf:
        return mag.amax(dim=axis, keepdim=True)
    return mag.pow(p).sum(dim=axis, keepdim=True).pow(1.0 / p)


def normalized_embeddings(X, p, value, axis):
    """Rescale every vector along `axis` to Lp norm `value` (no epsilon, like the reference)."""
    X = _as_tensor(X)
    return X / _lp_norm(X, p, axis) * value


def soft_constraint(X, p, value, axis):
    """sum(max(0, ||x||_p^p - value))."""
    return (_lp_norm(X, p, axis).pow(p) - value).clamp_min(0).sum()


def clip_constraint(X, p, value, axis):
    """Vectors whose Lp norm reaches `value` are rescaled onto it; shorter ones pass."""
    X = _as_tensor(X)
    nrm = _lp_norm(X, p, axis)
    return torch.where(nrm < value, X, X / nrm.clamp_min(1e-9) * value)


def Lp_regularization(X, p, axis):
    """sum(|x|^p) along `axis`."""
    return torch.abs(_as_tensor(X)).pow(p).sum(dim=axis)
