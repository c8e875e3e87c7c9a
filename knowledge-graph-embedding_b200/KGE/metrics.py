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
    """np.median's value (middle element, or the mean of the two middle ones) by one partition of the ranks,
    without its NaN scan: ranks are integers >= 1."""
    r = _r(ranks)
    n = r.size
    if n == 0:
        return float("nan")
    k = n // 2
    if n % 2:
        return float(np.partition(r, k)[k])
    p = np.partition(r, [k - 1, k])
    return float(0.5 * (p[k - 1] + p[k]))


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


def summary(ranks):
    """The nine metrics of KGEModel.evaluate (BaseModel.py:606-618 of the reference) in its key order, from one
    float64 copy of the ranks and one reciprocal — the same expressions as the functions above, so the same values."""
    r = _r(ranks)
    if r.size == 0:
        nan = float("nan")
        return {k: nan for k in ("mean_rank", "mean_reciprocal_rank", "median_rank", "geometric_mean_rank",
                                 "harmonic_mean_rank", "std_rank", "hit@1", "hit@3", "hit@10")}
    inv = 1.0 / r
    return {
        "mean_rank": float(np.mean(r)),
        "mean_reciprocal_rank": float(np.mean(inv)),
        "median_rank": median_rank(r),
        "geometric_mean_rank": float(np.exp(np.mean(np.log(r)))),
        "harmonic_mean_rank": float(r.size / np.sum(inv)),
        "std_rank": float(np.std(r)),
        "hit@1": float(np.mean(r <= 1)),
        "hit@3": float(np.mean(r <= 3)),
        "hit@10": float(np.mean(r <= 10)),
    }
