"""Counter-based negative sampling — numpy restatement (TEST INFRASTRUCTURE).

The reference draws negatives from *unseeded* global RNG state
(``KGE/ns_strategy.py:57-62`` ``tf.random.uniform``; ``KGE/utils.py:11-16``
``np.random.choice``), so "bit-exact negatives" can only be defined against our own
stream specification.  This module is that specification, written
independently of the CUDA code (numpy uint64 arithmetic, no shared source):

  generator : Philox4x32-10 (Salmon et al., SC'11), Random123 constants
  key       : (seed & 0xffffffff, seed >> 32)
  counter   : (n & 0xffffffff, n >> 32, step & 0xffffffff,
               ((step >> 32) << 1 | side_bit) & 0xffffffff)
              n = flat index i*k_side + j into the per-side sample array the
              reference sampler returns (shape (B*k_side,), ns_strategy.py:39,94),
              side_bit = 0 for 'h', 1 for 't'
  draw      : u = word 0 of the 4x32 output block
  uniform   : id = pool[u % len(pool)]          (pool=None means arange(N));
              TF's integer uniform is also lo + u % range.  No rejection of the
              true entity or of known positives (ns_strategy.py:57-64).
  typed     : P = members of type(ind2type[x]) in ascending id order, x the entity
              being replaced; if x in P at position q:  j = u % (len(P)-1),
              id = P[j + (j >= q)]   (== np.delete(P, q)[j], utils.py:12-14);
              if x not in P: id = P[u % len(P)].  An empty pool raises, as
              np.random.choice does in the reference.
"""
import numpy as np

_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = 0x9E3779B9
_W1 = 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)
_S32 = np.uint64(32)


def philox4x32_10(ctr, key):
    """ctr: (..., 4) uint32, key: (..., 2) uint32 (broadcastable) -> (..., 4) uint32."""
    ctr = np.asarray(ctr, dtype=np.uint64)
    key = np.asarray(key, dtype=np.uint64)
    c0, c1, c2, c3 = (ctr[..., i] for i in range(4))
    k0, k1 = key[..., 0], key[..., 1]
    for rnd in range(10):
        if rnd:
            k0 = (k0 + np.uint64(_W0)) & _MASK
            k1 = (k1 + np.uint64(_W1)) & _MASK
        p0 = _M0 * c0          # < 2**64, exact in uint64
        p1 = _M1 * c2
        hi0, lo0 = p0 >> _S32, p0 & _MASK
        hi1, lo1 = p1 >> _S32, p1 & _MASK
        c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
    return np.stack([c0, c1, c2, c3], axis=-1).astype(np.uint32)


def draw_u32(seed, step, side_bit, count):
    """u[n] for n in [0, count): word 0 of Philox at the counter defined above."""
    n = np.arange(count, dtype=np.uint64)
    ctr = np.empty((count, 4), dtype=np.uint64)
    ctr[:, 0] = n & _MASK
    ctr[:, 1] = n >> _S32
    ctr[:, 2] = np.uint64(step & 0xFFFFFFFF)
    ctr[:, 3] = np.uint64((((step >> 32) << 1) | side_bit) & 0xFFFFFFFF)
    key = np.array([seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF], dtype=np.uint64)
    return philox4x32_10(ctr, key)[:, 0]


def epoch_permutation(seed, epoch, n):
    """Order of pass `epoch` over n training triples (kge_epoch_perm, include/kge_abi.h): stable argsort of
    u[i] = draw_u32(seed, epoch, 1, n).  The reference shuffles with tf.data (data_utils.py:186-190, unseeded
    stream); the permutation is defined on the library's own counter-based stream like the negatives."""
    return np.argsort(draw_u32(seed, epoch, 1, n), kind="stable").astype(np.uint32)


def sample_uniform(seed, step, side, B, k_side, n_ent, pool=None):
    """Restates UniformStrategy.__call__ (ns_strategy.py:39-64) on our stream."""
    u = draw_u32(seed, step, 0 if side == "h" else 1, B * k_side).astype(np.int64)
    if pool is None:
        return u % n_ent
    pool = np.asarray(pool, dtype=np.int64)
    return pool[u % len(pool)]


class TypedPools:
    """CSR form of metadata['type2inds'] / ['ind2type'] (BaseModel.py:258-276)."""

    def __init__(self, ind2type=None, type2inds=None):
        if type2inds is None:
            ind2type = list(ind2type)
            names = sorted(set(ind2type), key=lambda s: str(s))
            type2inds = {t: np.array([i for i, ti in enumerate(ind2type) if ti == t],
                                     dtype=np.int64) for t in names}
        self.names = list(type2inds.keys())
        tid = {t: i for i, t in enumerate(self.names)}
        self.ent_type = np.array([tid[t] for t in ind2type], dtype=np.int64)
        sizes = [len(type2inds[t]) for t in self.names]
        self.offsets = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
        self.members = (np.concatenate([np.asarray(type2inds[t], dtype=np.int64)
                                        for t in self.names])
                        if self.names else np.zeros(0, np.int64))
        # position of each entity inside its own type pool, -1 when absent
        self.pos = np.full(len(self.ent_type), -1, dtype=np.int64)
        for ti in range(len(self.names)):
            seg = self.members[self.offsets[ti]:self.offsets[ti + 1]]
            for q, e in enumerate(seg):
                if e < len(self.ent_type) and self.ent_type[e] == ti and self.pos[e] < 0:
                    self.pos[e] = q


def sample_typed(seed, step, side, X, k_side, pools):
    """Restates TypedStrategy.__call__ + ns_with_same_type (ns_strategy.py:94-132,
    utils.py:11-16) on our stream."""
    X = np.asarray(X)
    ref = X[:, 0] if side == "h" else X[:, 2]
    B = len(ref)
    u = draw_u32(seed, step, 0 if side == "h" else 1, B * k_side).astype(np.int64)
    x = np.repeat(ref, k_side)
    t = pools.ent_type[x]
    lo = pools.offsets[t]
    ln = pools.offsets[t + 1] - lo
    q = pools.pos[x]
    eff = np.where(q >= 0, ln - 1, ln)
    if np.any(eff <= 0):
        raise ValueError("a cannot be empty unless no samples are taken")
    j = u % eff
    j = j + ((q >= 0) & (j >= q))
    return pools.members[lo + j]


def negatives(seed, step, X, negative_ratio, corrupt_side, n_ent, pool=None, typed=None):
    """Negative entity ids in the reference's row order (BaseModel.py:332-408).

    Returns (ids, k_eff): ids has shape (B*k_eff,), row i*k_eff + c belongs to
    positive i; for 'h+t' even c are head corruptions and odd c tail corruptions
    (the concat/reshape of BaseModel.py:354-356), k_eff = 2*(k//2).
    """
    X = np.asarray(X)
    B = len(X)

    def one(side, k):
        if typed is not None:
            return sample_typed(seed, step, side, X, k, typed)
        return sample_uniform(seed, step, side, B, k, n_ent, pool)

    if corrupt_side in ("h", "t"):
        return one(corrupt_side, negative_ratio), negative_ratio
    k2 = negative_ratio // 2
    nh = one("h", k2).reshape(B, k2)
    nt = one("t", k2).reshape(B, k2)
    return np.stack([nh, nt], axis=-1).reshape(-1), 2 * k2
