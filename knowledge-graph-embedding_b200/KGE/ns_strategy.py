"""Negative samplers — drop-in for KGE/ns_strategy.py of the reference
(UniformStrategy :23-64, TypedStrategy :67-132, worker KGE/utils.py:11-16).

Both are descriptors for the CUDA sampler: inside `train()` negatives are drawn in the
fused step kernel; called directly (`strategy(X, negative_ratio, side)`, as
tests/test_ns_strategy.py does) they launch `kge_sample_negatives` through the C-ABI.
The stream is counter-based Philox4x32-10 keyed by (seed, step, side, element) — the
reference's `tf.random.uniform` / `np.random.choice` are unseeded global streams
(SURVEY.md F9), so the *distribution* is what is reproduced:
  Uniform:  pool[u % len(pool)]                         (no rejection of true triples)
  Typed:    same-type pool of the replaced entity minus that entity, with replacement.
"""
import ctypes as C
import os

import numpy as np
import torch

from ._b200 import lib as L


def _type_csr(ind2type, type2inds=None):
    """metadata['ind2type'] (+ optional 'type2inds') -> CSR arrays for the device sampler."""
    labels = list(ind2type)
    if type2inds is None:
        names = sorted(set(labels), key=str)
        type2inds = {t: np.array([i for i, ti in enumerate(labels) if ti == t], dtype=np.int64) for t in names}
    names = list(type2inds.keys())
    tid = {t: i for i, t in enumerate(names)}
    ent_type = np.array([tid[t] for t in labels], dtype=np.int64)
    sizes = [len(type2inds[t]) for t in names]
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    members = np.concatenate([np.asarray(type2inds[t], dtype=np.int64) for t in names]) if names \
        else np.zeros(0, np.int64)
    pos = np.full(len(labels), -1, dtype=np.int64)
    for ti in range(len(names)):
        seg = members[off[ti]:off[ti + 1]]
        own = np.flatnonzero(ent_type[seg] == ti)
        # first position of each entity inside its own pool
        for q in own[::-1]:
            pos[seg[q]] = q
    return ent_type, off, members, pos


class NegativeSampler:
    def __init__(self):
        raise NotImplementedError("a NegativeSampler subclass must define its parameters in __init__()")

    def __call__(self, X, negative_ratio, side):
        raise NotImplementedError("a NegativeSampler subclass must implement __call__(X, negative_ratio, side)")

    # standalone stream state (train() uses its own seed / step counter)
    seed = None
    _step = 0

    def _next_seed_step(self):
        if self.seed is None:
            self.seed = int.from_bytes(os.urandom(7), "little")
        s = self._step
        self._step += 1
        return self.seed, s

    def _launch(self, X, negative_ratio, side, n_ent, fill_sampler):
        from ._b200 import engine as E
        if side not in ("h", "t"):
            raise ValueError("side must be 'h' or 't'")
        out_dtype = X.dtype if torch.is_tensor(X) else torch.as_tensor(np.asarray(X)).dtype
        dev = torch.device(os.environ.get("KGE_DEVICE", "cuda:0"))
        ids = torch.as_tensor(np.asarray(X) if not torch.is_tensor(X) else X).to(device=dev, dtype=torch.int64)
        ids = ids.contiguous()
        cfg = L.Cfg(abi_version=L.KGE_ABI_VERSION, model=L.TRANSE, d_ent=1, d_rel=1, n_ent=int(n_ent), n_rel=1,
                    id_dtype=1, score_p=2.0, negative_ratio=int(negative_ratio),
                    corrupt_side=L.SIDE_H if side == "h" else L.SIDE_T)
        smp, keep = fill_sampler(cfg, dev)
        out = torch.empty(ids.shape[0] * int(negative_ratio), dtype=torch.int64, device=dev)
        seed, step = self._next_seed_step()
        L.check(L.load().kge_sample_negatives(E._ctx(dev.index or 0), C.byref(cfg), C.byref(smp),
                                              C.c_void_p(ids.data_ptr()), ids.shape[0], seed, step,
                                              C.c_void_p(out.data_ptr()), E._stream()))
        del keep
        return out.to(out_dtype)


class UniformStrategy(NegativeSampler):
    """Every entity of `sample_pool` is equally likely."""

    def __init__(self, sample_pool):
        self.sample_pool = sample_pool

    def pool_array(self):
        p = self.sample_pool
        if torch.is_tensor(p):
            p = p.detach().cpu().numpy()
        return np.asarray(p, dtype=np.int64)

    def __call__(self, X, negative_ratio, side):
        pool = self.pool_array()

        def fill(cfg, dev):
            cfg.sampler = L.NS_UNIFORM
            t = torch.as_tensor(pool).to(dev)
            return L.Sampler(pool=t.data_ptr(), pool_len=t.numel()), [t]
        return self._launch(X, negative_ratio, side, int(pool.max()) + 1 if pool.size else 1, fill)


class TypedStrategy(NegativeSampler):
    """Replace an entity only by another entity of the same type (never by itself).
    `pool` (a multiprocessing pool in the reference) is accepted and ignored: sampling is
    one kernel launch."""

    def __init__(self, pool, metadata):
        self.pool = pool
        self.metadata = metadata

    def csr(self):
        return _type_csr(self.metadata["ind2type"], self.metadata.get("type2inds"))

    def __call__(self, X, negative_ratio, side):
        ent_type, off, members, pos = self.csr()
        check_typed_pools(ent_type, off, pos)

        def fill(cfg, dev):
            cfg.sampler = L.NS_TYPED
            ts = [torch.as_tensor(a).to(dev) for a in (ent_type, off, members, pos)]
            return L.Sampler(ent_type=ts[0].data_ptr(), type_off=ts[1].data_ptr(), type_members=ts[2].data_ptr(),
                             ent_pos=ts[3].data_ptr(), n_types=len(off) - 1), ts
        return self._launch(X, negative_ratio, side, len(ent_type), fill)


def check_typed_pools(ent_type, off, pos):
    """A type with a single member leaves an empty pool once the entity itself is removed;
    the reference fails there inside np.random.choice (KGE/utils.py:13-14)."""
    sizes = off[1:] - off[:-1]
    eff = sizes[ent_type] - (pos >= 0)
    if np.any(eff <= 0):
        raise ValueError("'a' cannot be empty unless no samples are taken: entity type with no other member")
