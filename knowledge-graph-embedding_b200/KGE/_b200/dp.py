"""Host side of multi-GPU runs: who owns which positives / entities, and the
one collective of a data-parallel step.  torch.distributed is the plumbing (NCCL on GPUs; the
same code runs over gloo on CPU tensors in tests/test_dp_gloo.py).

  * evaluation: every rank scores all queries against its own slice of the entity table
    (`entity_shard`) and the int32 "better candidate" counts are summed (`allreduce_counts`);
  * training: every rank holds all tables and `batch_shard` of the step's positives; the
    per-table summed gradients and the five scalars (norm^2 x 4, loss) live in ONE flat fp32
    buffer (`GradBuffer`) so a step needs a single all-reduce; each rank then applies the
    identical update, so replicas stay bit-identical;
  * on one NVLink / NVSwitch node the all-reduce is replaced by `PeerStaging`: the gradient
    kernels store their row sums straight into every rank's staging slot (peer memory), one
    cross-GPU barrier follows, and the update kernel sums the slots in rank order.
"""
import os

import torch
import torch.distributed as dist


def world():
    """(rank, world_size) of the default process group, (0, 1) when not initialised."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def split_even(n, parts, i):
    """[lo, hi) of part i when n items are split into `parts` contiguous near-equal ranges
    (the first n % parts ranges get one more)."""
    if not (0 <= i < parts):
        raise ValueError(f"part {i} out of range [0, {parts})")
    q, r = divmod(int(n), int(parts))
    lo = i * q + min(i, r)
    return lo, lo + q + (1 if i < r else 0)


def entity_shard(n_ent, rank=None, world_size=None):
    r, w = world()
    return split_even(n_ent, w if world_size is None else world_size, r if rank is None else rank)


def batch_shard(B_global, rank=None, world_size=None):
    """Positives [i_base, i_base + B) of a global batch owned by `rank`."""
    r, w = world()
    return split_even(B_global, w if world_size is None else world_size, r if rank is None else rank)


def allreduce_counts(counts, group=None):
    """Sum of the per-shard integer counts; exact and order independent."""
    if world()[1] > 1:
        dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=group)
    return counts


class GradBuffer:
    """One flat fp32 buffer [table 0 | table 1 | table 2 | table 3 | scal(8)].
    `tables` = {slot: (rows, ld)}; `view(slot)` is that table's (rows, ld) window and `scal`
    the trailing 8 floats (norm^2 of slots 0..3, loss share, 3 unused)."""

    N_SCAL = 8

    def __init__(self, tables, device):
        self.offsets = {}
        o = 0
        for s in sorted(tables):
            rows, ld = tables[s]
            if ld % 4:
                raise ValueError("row stride must be a multiple of 4 floats")
            self.offsets[s] = (o, rows, ld)
            o += rows * ld
        self.scal_off = o
        self.flat = torch.zeros(o + self.N_SCAL, dtype=torch.float32, device=device)

    def view(self, slot):
        o, rows, ld = self.offsets[slot]
        return self.flat[o:o + rows * ld].view(rows, ld)

    @property
    def scal(self):
        return self.flat[self.scal_off:self.scal_off + self.N_SCAL]

    def allreduce(self, group=None):
        """The step's single collective.  NCCL / gloo sum in a rank-independent order, so
        every rank ends with the same bits."""
        if world()[1] > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)
        return self


class PeerTables:
    """Weight tables (and the entity table's shadow `wn`) of one rank in peer-mapped memory, for the
    owner-partitioned exchange (kge_grads.owner_rows): the owner of a row stores the updated row into
    every rank's table.  Layout [slot a: w | slot a: wn | slot b: w | ...] in floats, 256-byte aligned
    sections; `tables` = {slot: (rows, ld)}, `wn_slots` = slots that carry a shadow table."""

    def __init__(self, tables, wn_slots=()):
        self.off, o = {}, 0
        for s in sorted(tables):
            rows, ld = tables[s]
            for kind in ("w", "wn") if s in wn_slots else ("w",):
                self.off[(kind, s)] = (o, rows, ld)
                o += (rows * ld + 63) // 64 * 64
        self.numel = o
        self.buf, self.base_ptrs, self._keep, self.mc_ptr = None, None, None, 0

    def view(self, kind, s):
        o, rows, ld = self.off[(kind, s)]
        return self.buf[o:o + rows * ld].view(rows, ld)

    def addr(self, owner, kind, s):
        return self.base_ptrs[owner] + 4 * self.off[(kind, s)][0]

    @classmethod
    def local(cls, tables, wn_slots, world_size, device):
        """Emulated ranks that share ONE replica (tests): every 'peer' is the same buffer."""
        t = cls(tables, wn_slots)
        t.buf = torch.zeros(max(t.numel, 1), dtype=torch.float32, device=device)
        t.base_ptrs = [t.buf.data_ptr()] * world_size
        return t

    @classmethod
    def symmetric(cls, tables, wn_slots, device, group=None):
        import torch.distributed._symmetric_memory as sm
        t = cls(tables, wn_slots)
        t.buf = sm.empty(max(t.numel, 1), dtype=torch.float32, device=device)
        hdl = sm.rendezvous(t.buf, group=group if group is not None else dist.group.WORLD)
        t.buf.zero_()
        t.base_ptrs = [int(p) for p in hdl.buffer_ptrs]
        # NVSwitch multicast mapping of the same buffer (0 where NVLS is not available): one store, every replica
        t.mc_ptr = 0
        if os.environ.get("KGE_DP_MULTICAST", "1") != "0":
            try:
                t.mc_ptr = int(getattr(hdl, "multicast_ptr", 0) or 0)
            except Exception:                                       # noqa: BLE001
                t.mc_ptr = 0
        t._keep = hdl
        return t


class PeerStaging:
    """Staging memory of the fused exchange, identical on every rank:
    [2 step parities][world slots][table 0 | .. | table 3 | scal(8)] fp32.
    `base_ptrs[j]` = address of rank j's staging buffer as mapped into THIS process
    (torch symmetric memory on a real node; plain local tensors when ranks are emulated on one
    device in tests).  `push(parity)` / `slots(parity)` fill the `peer` fields of kge_grads for
    kge_train_grad / kge_train_apply (include/kge_abi.h)."""

    def __init__(self, tables, world_size, rank, base_ptrs=None, barrier=None, keep=None, ptabs=None, w_private=False):
        from . import lib as L
        if not (1 <= world_size <= L.KGE_MAX_PEERS):
            raise ValueError(f"peer exchange supports up to {L.KGE_MAX_PEERS} ranks")
        self.world, self.rank = int(world_size), int(rank)
        self.offsets, o = {}, 0
        for s in sorted(tables):
            rows, ld = tables[s]
            if ld % 4:
                raise ValueError("row stride must be a multiple of 4 floats")
            self.offsets[s] = o
            o += rows * ld
        self.scal_off = o
        self.per = (o + GradBuffer.N_SCAL + 63) // 64 * 64          # slots stay 256-byte aligned
        self.base_ptrs = list(base_ptrs) if base_ptrs is not None else None
        self.barrier = barrier
        self._keep = keep                                           # owners of the memory
        self._cache = {}
        # owner-partitioned exchange: rank j owns rows [j * owner_rows[s], (j+1) * owner_rows[s]) of table s
        self.ptabs = ptabs
        self.w_private = bool(w_private)      # kge_grads.w_private: only wn is sent to the replicas
        self.owner_rows = {s: -(-tables[s][0] // self.world) for s in tables}

    def numel(self):
        return 2 * self.world * self.per + 64          # + the flag words of kge_peer_barrier

    def peer_flags(self):
        """kge_peer_flags over the last 64 words of every rank's staging buffer (zeroed at set-up)."""
        from . import lib as L
        if "flags" not in self._cache:
            F = L.PeerFlags(world=self.world, rank=self.rank)
            for j in range(self.world):
                F.flags[j] = self.base_ptrs[j] + 4 * (2 * self.world * self.per)
            self._cache["flags"] = F
        return self._cache["flags"]

    def check_alive(self):
        """Raise if an in-stream barrier (kge_peer_barrier) ever timed out on this rank: the error word
        flags[rank][KGE_MAX_PEERS] is sticky and the apply halves have been skipping their updates since.
        Costs a device sync: call where the host synchronises anyway (per epoch)."""
        from . import lib as L
        buf = self._keep[0] if isinstance(self._keep, tuple) else self._keep
        if buf is None or not torch.is_tensor(buf):
            return
        word = buf[2 * self.world * self.per + L.KGE_MAX_PEERS:2 * self.world * self.per + L.KGE_MAX_PEERS + 1]
        if int(word.view(torch.int32).item()) == L.PEER_DEAD:
            raise RuntimeError("data-parallel step: a peer GPU did not reach the in-stream barrier within ~20 s "
                               "(kge_peer_barrier gave up); the tables of this rank were left at the last completed step")

    def _addr(self, owner, parity, slot, off):
        return self.base_ptrs[owner] + 4 * ((parity * self.world + slot) * self.per + off)

    def _grads(self, parity, push):
        from . import lib as L
        key = (parity, push)
        if key not in self._cache:
            G = L.Grads()
            G.n_peers = self.world
            for j in range(self.world):
                owner, slot = (j, self.rank) if push else (self.rank, j)
                for s, off in self.offsets.items():
                    G.peer[j][s] = self._addr(owner, parity, slot, off)
                G.peer_scal[j] = self._addr(owner, parity, slot, self.scal_off)
            if not push:            # the apply half skips its update once the barrier's error word is set
                G.abort_word = self.base_ptrs[self.rank] + 4 * (2 * self.world * self.per + L.KGE_MAX_PEERS)
            if self.ptabs is not None:
                G.owner_rank = self.rank
                G.w_private = int(self.w_private)
                for s in self.offsets:
                    G.owner_rows[s] = self.owner_rows[s]
                    for j in range(self.world):
                        G.w_peer[j][s] = self.ptabs.addr(j, "w", s)
                        if ("wn", s) in self.ptabs.off:
                            G.wn_peer[j] = self.ptabs.addr(j, "wn", s)
                    if self.ptabs.mc_ptr and not push:
                        G.w_mc[s] = self.ptabs.mc_ptr + 4 * self.ptabs.off[("w", s)][0]
                        if ("wn", s) in self.ptabs.off:
                            G.wn_mc = self.ptabs.mc_ptr + 4 * self.ptabs.off[("wn", s)][0]
            self._cache[key] = G
        return self._cache[key]

    def push(self, parity):
        """Where this rank's gradients go on every rank."""
        return self._grads(parity & 1, True)

    def slots(self, parity):
        """Where every rank's gradients sit in this rank's memory."""
        return self._grads(parity & 1, False)

    @classmethod
    def symmetric(cls, tables, device, group=None, wn_slots=None, w_private=False):
        """Staging in torch symmetric memory (CUDA VMM peer mappings over NVLink), or None when
        any rank cannot set it up (the caller then keeps the NCCL all-reduce).  wn_slots not None:
        owner-partitioned exchange — the weight tables get peer-mapped memory too (`ptabs`)."""
        rank, world_size = world()
        ok, st = 1, None
        try:
            import torch.distributed._symmetric_memory as sm
            st = cls(tables, world_size, rank, w_private=w_private and bool(wn_slots))
            if wn_slots is not None:
                st.ptabs = PeerTables.symmetric(tables, wn_slots, device, group)
            buf = sm.empty(st.numel(), dtype=torch.float32, device=device)
            hdl = sm.rendezvous(buf, group=group if group is not None else dist.group.WORLD)
            buf.zero_()
            st.base_ptrs = [int(p) for p in hdl.buffer_ptrs]
            st.barrier = hdl.barrier
            st._keep = (buf, hdl)
        except Exception as e:                                      # noqa: BLE001 - any failure => NCCL path
            ok = 0
            if os.environ.get("KGE_DP_EXCHANGE") == "peer":
                raise RuntimeError(f"KGE_DP_EXCHANGE=peer but symmetric memory is unavailable: {e!r}")
        flag = torch.tensor([ok], dtype=torch.int32, device=device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        if int(flag.item()) != 1:
            return None
        hdl.barrier()
        return st

    def sync_weights(self, tabs, group=None):
        """w_private: every owner broadcasts its rows of w of the shadow-carrying tables (collective)."""
        if not (self.w_private and self.ptabs is not None):
            return
        for s, pt in tabs.items():
            if ("wn", s) not in self.ptabs.off:
                continue
            n = self.owner_rows[s]
            for j in range(self.world):
                lo, hi = j * n, min((j + 1) * n, pt.rows)
                if hi > lo:
                    dist.broadcast(pt.storage[lo:hi], src=j, group=group)


def exchange_kind(world_size, dense_regulariser=False):
    """'owner' (reduce-scatter / all-gather carried by kernel stores into peer memory: row sums go to
    the row's owner, the owner updates and stores the new row into every replica; two barriers),
    'peer' (row sums stored into every rank's staging, every rank updates everything; one barrier)
    or 'nccl' (all-reduce of the flat buffer).  Default: owner wherever it applies (per-rank NVLink
    bytes and update work do not grow with the number of ranks); KGE_DP_EXCHANGE=owner|peer|nccl
    overrides."""
    e = os.environ.get("KGE_DP_EXCHANGE", "")
    if e in ("peer", "nccl") or (e == "owner" and not dense_regulariser):
        return e
    if world_size < 2:
        return "nccl"
    if not dense_regulariser:
        return "owner"
    return "peer" if world_size <= 4 else "nccl"
