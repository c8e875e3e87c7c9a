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


class ShardedTrainer:
    """What KGEModel.train / evaluate drive when the entity tables are sharded: either
    this process's rank of a torchrun job (`emulate` = 0: one ShardedEngine, peers over torch symmetric
    memory, collectives over the default process group) or `emulate` = G ranks on ONE device (tests,
    single-GPU dry runs: plain buffers, phases in lock-step, the all-reduce replaced by a rank-ordered sum).

    `values` = {slot: full (rows, ...) tensor}: every rank is given the same full initial tables (same
    seed) and keeps its own rows of the entity slots."""

    def __init__(self, cfg_kwargs, values, B_global_max, device, emulate=0, group=None):
        self.device, self.group = torch.device(device), group
        self.emulated = int(emulate) > 1
        self.rank, self.world = (0, int(emulate)) if self.emulated else _dp.world()
        self.cfg_kwargs = dict(cfg_kwargs)
        self.n_ent = int(cfg_kwargs["n_ent"])
        self.S = shard_rows(self.n_ent, self.world)
        ent = {s: torch.as_tensor(v, dtype=torch.float32) for s, v in values.items() if s < 2}
        rel = {s: v for s, v in values.items() if s >= 2}
        self.shapes = {s: tuple(v.shape) for s, v in ent.items()}
        lds = {s: (int(np.prod(v.shape[1:])) + 3) // 4 * 4 for s, v in ent.items()}
        probe = _engine.Engine.__new__(_engine.Engine)          # k_eff without building an engine
        probe.cfg = L.Cfg(negative_ratio=int(cfg_kwargs["negative_ratio"]), corrupt_side=int(cfg_kwargs["corrupt_side"]))
        k_eff = probe.k_eff()
        self.B_local_max = -(-int(B_global_max) // self.world)
        cap = (2 + k_eff) * self.B_local_max
        rows = lambda g: (g * self.S, min(self.n_ent, (g + 1) * self.S))      # noqa: E731
        if self.emulated:
            mems = ShardMemory.local(self.S, lds, self.world, cap, self.device)
            self.ranks = [ShardedEngine(cfg_kwargs, {s: v[rows(g)[0]:rows(g)[1]] for s, v in ent.items()}, rel, g,
                                        self.world, mems[g], self.B_local_max, device=self.device)
                          for g in range(self.world)]
        else:
            mem = ShardMemory.symmetric(self.S, lds, cap, self.device, group)
            lo, hi = rows(self.rank)
            self.ranks = [ShardedEngine(cfg_kwargs, {s: v[lo:hi] for s, v in ent.items()}, rel, self.rank, self.world,
                                        mem, self.B_local_max, device=self.device)]
        self._eval_engine = None

    # ---- engines ------------------------------------------------------------------------------
    @property
    def engines(self):
        return [r.eng for r in self.ranks]

    def each_engine(self, fn):
        for r in self.ranks:
            fn(r.eng)

    # ---- training -----------------------------------------------------------------------------
    def train_step(self, batch, seed, step, loss_out=None):
        """One step on the GLOBAL batch (identical on every rank): each rank takes its contiguous share."""
        B = int(batch.shape[0])
        self._eval_engine = None
        if self.emulated:
            parts, bases = [], []
            for g in range(self.world):
                lo, hi = _dp.split_even(B, self.world, g)
                parts.append(batch[lo:hi])
                bases.append(lo)
            if min(len(p) for p in parts) == 0:
                raise ValueError(f"batch of {B} positives cannot be split over {self.world} ranks")
            loss = emulate_step(self.ranks, parts, seed, step, B, bases)
            if loss_out is not None:
                loss_out.copy_(loss)
                return loss_out
            return loss
        lo, hi = _dp.split_even(B, self.world, self.rank)
        if hi == lo:
            raise ValueError(f"batch of {B} positives cannot be split over {self.world} ranks")
        return self.ranks[0].train_step(batch[lo:hi], seed, step, B, lo, loss_out=loss_out, group=self.group)

    # ---- reading / writing whole tables -----------------------------------------------------------
    def full_table(self, slot):
        """(n_ent, ...) entity table or the replicated relation table, materialised (flushed), on this rank."""
        if slot >= 2:
            return self.ranks[0].eng.tab[slot].view
        if self.emulated:
            flat = torch.cat([r.local_rows(slot).reshape(r.hi - r.lo, -1) for r in self.ranks], 0)
        else:
            flat = self.ranks[0].gather_entity_table(slot, self.group)
        return flat.reshape(self.shapes[slot])

    def assign_table(self, slot, value):
        """Overwrite a table from a full tensor (restore / assign): every rank keeps its rows; optimizer state
        of the touched rows is kept, the lazy stamps are brought to the current step first."""
        value = torch.as_tensor(value, dtype=torch.float32).to(self.device)
        self._eval_engine = None
        for r in self.ranks:
            r.eng.flush()
            pt = r.eng.tab[slot]
            if slot >= 2:
                pt.view.copy_(value.reshape(pt.shape))
            else:
                pt.view.copy_(value[r.lo:r.hi].reshape(pt.shape))
            pt.touch()

    def eval_engine(self):
        """Engine on the gathered (replicated) tables for scoring / ranking: only the TRAINING state
        (w, m, v, last) is sharded; evaluation keeps w replicated (2.56 GB at C5) and shards the candidates
        (kge_rank ent_lo / ent_hi + all-reduce of the counts).  Cached until the next training step."""
        if self._eval_engine is None:
            tabs = {s: self.full_table(s) for s in self.ranks[0].eng.tab}
            self._eval_engine = _engine.Engine(self.cfg_kwargs, tabs, device=self.device)
            self._eval_engine.sampler, self._eval_engine._sampler_keep = self.ranks[0].eng.sampler, self.ranks[0].eng._sampler_keep
            self._eval_engine.cfg.sampler = self.ranks[0].eng.cfg.sampler
        return self._eval_engine
