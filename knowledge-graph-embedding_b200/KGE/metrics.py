"""Rank aggregations — drop-in for KGE/metrics.py of the reference (:5-25): trivial host
reductions over the integer ranks the ranking kernel produced."""
import numpy as np


def _r(ranks):
    return np.asarray(ranks, dtype=np.float64)


def mean_reciprocal_rank(ranks):
    return float(np.mean(1.0 / _r(ranks)))


def mean_rank(ranks):
    return float(np.mean(_r(ranks)))


def median_rank(ranks):
    return float(np.median(_r(ranks)))


def geometric_mean_rank(ranks):
    return float(np.exp(np.mean(np.log(_r(ranks)))))


def harmonic_mean_rank(ranks):
    r = _r(ranks)
    return float(r.size / np.sum(1.0 / r))


def std_rank(ranks):
    return float(np.std(_r(ranks)))


def hits_at_k(ranks, k):
    assert k >= 1, "k needs >= 1"
    return float(np.mean(_r(ranks) <= k))
