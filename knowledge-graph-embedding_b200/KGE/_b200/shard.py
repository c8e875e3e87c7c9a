"""Entity tables sharded over the GPUs of one NVLink / NVSwitch node: host side of
kge_shard_lookup / _pre / _grad / _apply (include/kge_abi.h, csrc/shard_kernels.cuh).

Rank g owns entity rows [g S, (g+1) S).  Its peer-visible memory is ONE buffer
    [ w slot 0 | w slot 1 | id inbox | count inbox | value inbox slot 0 | value inbox slot 1 ]
allocated from torch symmetric memory (CUDA VMM mappings over NVLink) so that every rank holds the
address of every other rank's buffer; m, v, `last` and the de-duplication scratch stay private.
In tests several ranks are emulated on one device with plain tensors (`ShardMemory.local`)."""
import ctypes as C

import numpy as np
import torch
import torch.distributed as dist

from . import dp as _dp, engine as _engine, lib as L


def shard_rows(n_ent, world_size):
    """Rows per shard: ceil(n_ent / world) rounded up to 64 rows."""
    s = -(-int(n_ent) // int(world_size))
    return (s + 63) // 64 * 64


class ShardMemory:
    """Layout (in floats / 4-byte words) of one rank's peer-visible buffer."""

    def __init__(self, S, lds, world_size, cap):
        self.S, self.lds, self.world, self.cap = int(S), dict(lds), int(world_size), int(cap)
        o = 0

        def take(n):
            nonlocal o
            at = o
            o += (int(n) + 63) // 64 * 64          # sections stay 256-byte aligned
            return at
        self.off_w = {s: take(self.S * ld) for s, ld in sorted(self.lds.items())}
        self.off_ids = take(self.world * self.cap)
        self.off_cnt = take(self.world)
        self.off_val = {s: take(self.world * self.cap * ld) for s, ld in sorted(self.lds.items())}
        self.numel = o
        self.buf = None
        self.base_ptrs = None
        self.barrier = lambda: None
        self._keep = None

    def addr(self, owner, off):
        return self.base_ptrs[owner] + 4 * off

    @classmethod
    def local(cls, S, lds, world_size, cap, device):
        """Emulated ranks on one device: `world_size` plain buffers, no barrier needed (one stream)."""
        mems = [cls(S, lds, world_size, cap) for _ in range(world_size)]
        for m in mems:
            m.buf = torch.zeros(m.numel, dtype=torch.float32, device=device)
        ptrs = [m.buf.data_ptr() for m in mems]
        for m in mems:
            m.base_ptrs = ptrs
        return mems

    @classmethod
    def symmetric(cls, S, lds, cap, device, group=None):
        """This process's buffer in torch symmetric memory; raises when peer mappings are unavailable
        (the sharded step has no other transport)."""
        import torch.distributed._symmetric_memory as sm
        rank, world_size = _dp.world()
        m = cls(S, lds, world_size, cap)
        m.buf = sm.empty(m.numel, dtype=torch.float32, device=device)
        hdl = sm.rendezvous(m.buf, group=group if group is not None else dist.group.WORLD)
        m.buf.zero_()
        m.base_ptrs = [int(p) for p in hdl.buffer_ptrs]
        m.barrier = hdl.barrier
        m._keep = hdl
        torch.cuda.synchronize()
        hdl.barrier()
        return m


class ShardedEngine:
    """One rank of the sharded step.  `ent_tables` = {slot: (S_local, cols) values of THIS rank's rows},
    `rel_tables` = {slot: full values} (replicated)."""

    def __init__(self, cfg_kwargs, ent_tables, rel_tables, rank, world_size, mem, B_local_max, device="cuda:0"):
        self.rank, self.world, self.mem = int(rank), int(world_size), mem
        self.device = torch.device(device)
        S = mem.S
        n_ent = int(cfg_kwargs["n_ent"])
        self.lo = self.rank * S
        self.hi = min(n_ent, self.lo + S)
        tabs = dict(rel_tables)
        tabs.update(ent_tables)
        self.eng = _engine.Engine(cfg_kwargs, tabs, device=device, lazy=True)
        # entity weights move into the peer-visible buffer
        for s in ent_tables:
            pt = self.eng.tab[s]
            assert pt.rows == self.hi - self.lo and pt.ld == mem.lds[s], (pt.rows, self.hi - self.lo, pt.ld, mem.lds[s])
            region = mem.buf[mem.off_w[s]:mem.off_w[s] + pt.rows * pt.ld].view(pt.rows, pt.ld)
            region.copy_(pt.storage)
            pt.storage = region
        k = self.eng.k_eff()
        need = (2 + k) * int(B_local_max)
        if mem.cap < need:
            raise ValueError(f"inbox capacity {mem.cap} < {need} looked-up rows of a local batch")
        n_work = min(S, self.world * mem.cap)
        z = lambda n: torch.zeros(int(n), dtype=torch.int32, device=self.device)   # noqa: E731
        self._scratch = [z(S), z(S), z(4 + 3 * n_work), z(n_work * self.world)]
        sh = L.Shard(world=self.world, rank=self.rank, shard_rows=S, cap=mem.cap)
        for j in range(self.world):
            for s in ent_tables:
                sh.w_peer[j][s] = mem.addr(j, mem.off_w[s])
                sh.val_peer[j][s] = mem.addr(j, mem.off_val[s])
            sh.ids_peer[j] = mem.addr(j, mem.off_ids)
            sh.cnt_peer[j] = mem.addr(j, mem.off_cnt)
        sh.stamp, sh.idx_of, sh.work, sh.contrib = (t.data_ptr() for t in self._scratch)
        self.sh = sh
        self.gbuf = _dp.GradBuffer({s: (pt.rows, pt.ld) for s, pt in self.eng.tab.items() if s >= 2}, self.device)
        self._G = L.Grads()
        for s in self.gbuf.offsets:
            self._G.dense[s] = self.gbuf.view(s).data_ptr()
        self._step_args = None

    # ---- the four calls ---------------------------------------------------------------------
    def lookup(self, triples_local, seed, step, B_global, i_base, neg_override=None):
        e = self.eng
        X = e._ids(triples_local)
        B = X.shape[0]
        ws = e._workspace(B, 0)
        dpd = L.Dp(world=self.world, rank=self.rank, B_global=B_global, i_base=i_base)
        neg = None
        if neg_override is not None:
            neg = torch.as_tensor(neg_override, dtype=torch.int64).to(self.device).contiguous()
        T = e._tables(with_opt=True)
        self._step_args = (X, B, ws, dpd, neg, seed, step)
        L.check(L.load().kge_shard_lookup(e.ctx, C.byref(e.cfg), C.byref(T), C.byref(e.sampler), C.byref(self.sh),
                                          C.byref(dpd), _engine._ptr(X), B, seed, step, _engine._ptr(neg),
                                          _engine._ptr(ws), ws.numel(), _engine._stream()))

    def pre(self):
        e = self.eng
        T = e._tables(with_opt=True)
        L.check(L.load().kge_shard_pre(e.ctx, C.byref(e.cfg), C.byref(T), C.byref(self.sh), e.adam_t + 1,
                                       _engine._stream()))

    def grad(self, want_negatives=False):
        e = self.eng
        X, B, ws, dpd, neg, seed, step = self._step_args
        neg_out = torch.empty(B * e.k_eff(), dtype=torch.int64, device=self.device) if want_negatives else None
        T = e._tables(with_opt=True)
        L.check(L.load().kge_shard_grad(e.ctx, C.byref(e.cfg), C.byref(T), C.byref(e.sampler), C.byref(self.sh),
                                        C.byref(dpd), _engine._ptr(X), B, seed, step, _engine._ptr(neg),
                                        C.byref(self._G), _engine._ptr(self.gbuf.scal), _engine._ptr(neg_out),
                                        _engine._ptr(ws), ws.numel(), _engine._stream()))
        return neg_out

    def apply(self, loss_out=None):
        e = self.eng
        X, B, ws, dpd, neg, seed, step = self._step_args
        if loss_out is None:
            loss_out = torch.empty(1, dtype=torch.float32, device=self.device)
        e.adam_t += 1
        T = e._tables(with_opt=True)
        L.check(L.load().kge_shard_apply(e.ctx, C.byref(e.cfg), C.byref(T), C.byref(self.sh), C.byref(self._G),
                                         _engine._ptr(self.gbuf.scal), B, e.adam_t, _engine._ptr(loss_out),
                                         _engine._ptr(ws), ws.numel(), _engine._stream()))
        e._sync_shadow(T)
        e._lazy_dirty = True
        return loss_out

    def train_step(self, triples_local, seed, step, B_global, i_base, loss_out=None, group=None):
        """One step in a real multi-process run; every rank must call it."""
        self.lookup(triples_local, seed, step, B_global, i_base)
        self.mem.barrier()
        self.pre()
        self.mem.barrier()
        self.grad()
        self.gbuf.allreduce(group)          # relation gradients, norms, loss — and the barrier before apply
        return self.apply(loss_out)

    # ---- reading the model ----------------------------------------------------------------------
    def local_rows(self, slot=0):
        """This rank's rows of an entity table, materialised (flushed)."""
        return self.eng.tab[slot].view

    def gather_entity_table(self, slot=0, group=None):
        """Full (n_ent, cols) entity table on every rank (evaluation keeps w replicated: 2.56 GB at C5;
        only the training state w + m + v + last is sharded)."""
        pt = self.eng.tab[slot]
        loc = torch.zeros((self.mem.S, pt.cols), dtype=torch.float32, device=self.device)
        loc[:pt.rows] = pt.view
        if self.world == 1 or not (dist.is_available() and dist.is_initialized()):
            return loc[:pt.rows]
        full = torch.empty((self.world * self.mem.S, pt.cols), dtype=torch.float32, device=self.device)
        dist.all_gather_into_tensor(full, loc, group=group)
        return full[:int(self.eng.cfg.n_ent)]


def emulate_step(ranks, batches, seed, step, B_global, bases, neg_overrides=None):
    """Drive one step of emulated ranks (one process, one stream): phases in lock-step, the all-reduce
    replaced by a rank-ordered sum of the flat buffers.  Returns the loss tensor of rank 0."""
    for r, X, b in zip(ranks, batches, bases):
        r.lookup(X, seed, step, B_global, b, None if neg_overrides is None else neg_overrides[r.rank])
    for r in ranks:
        r.pre()
    for r in ranks:
        r.grad()
    total = ranks[0].gbuf.flat.clone()
    for r in ranks[1:]:
        total += r.gbuf.flat
    for r in ranks:
        r.gbuf.flat.copy_(total)
    return [r.apply() for r in ranks][0]
