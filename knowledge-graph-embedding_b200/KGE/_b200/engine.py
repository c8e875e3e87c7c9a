"""Device-side state of one model and the calls into libkge_b200.so.

torch is plumbing only: it owns device memory (tables, Adam slots, workspace) and the
stream; every computation on the hot path is a kernel of the library.
"""
import ctypes as C
import os
import math
import numpy as np
import torch
from . import lib as L

_CTX = {}


def _ctx(device_index):
    if device_index not in _CTX:
        if not torch.cuda.is_available():
            raise RuntimeError("KGE (B200 build) needs a CUDA device: there is no CPU fallback")
        h = C.c_void_p()
        L.check(L.load().kge_ctx_create(device_index, C.byref(h)))
        if os.environ.get("KGE_RANK_TC_PASSES") is not None:  # A/B switch: MMAs per K step of the ranking GEMM (3 | 2)
            L.check(L.load().kge_set_option(h, 6, int(os.environ["KGE_RANK_TC_PASSES"])))
        if os.environ.get("KGE_PRE_EARLY") is not None:          # A/B switch: look-ahead grouping beside the step kernel
            L.check(L.load().kge_set_option(h, 8, int(os.environ["KGE_PRE_EARLY"])))
        if os.environ.get("KGE_CLUSTER_GROUP") is not None:      # A/B switch: one-kernel cluster grouping
            L.check(L.load().kge_set_option(h, 7, int(os.environ["KGE_CLUSTER_GROUP"])))
        if os.environ.get("KGE_RING_SLOTS") is not None:      # A/B switch: depth of the step kernel's TMA row ring
            L.check(L.load().kge_set_option(h, 4, int(os.environ["KGE_RING_SLOTS"])))
        _CTX[device_index] = h
    return _CTX[device_index]


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class _Staging:
    """Pinned staging buffer per device for large pageable host arrays (numpy inputs of evaluate(): positive_X is
    11.6 MB at FB15k size): a threaded host copy into pinned memory + ONE asynchronous DMA instead of the driver's
    synchronous pageable path.  The buffer is reused; an event guards it until the DMA that reads it has finished."""
    by_device = {}

    def __init__(self):
        self.buf, self.ev = None, None


def upload(t_cpu, device, dtype=None):
    """CPU tensor -> tensor on `device` (optionally cast there), asynchronously when the source is pinned or large
    enough to go through the staging buffer."""
    device = torch.device(device)
    nbytes = t_cpu.numel() * t_cpu.element_size()
    if (t_cpu.is_cuda or device.type != "cuda" or nbytes < (1 << 20) or nbytes > (1 << 30) or t_cpu.is_pinned()
            or not t_cpu.is_contiguous()):
        out = t_cpu.to(device=device, non_blocking=True)
    else:
        st = _Staging.by_device.setdefault(device, _Staging())
        if st.ev is not None:
            st.ev.synchronize()
        if st.buf is None or st.buf.numel() < nbytes:
            st.buf = torch.empty(1 << max(20, (nbytes - 1).bit_length()), dtype=torch.uint8, pin_memory=True)
        view = st.buf[:nbytes].view(t_cpu.dtype).reshape(t_cpu.shape)
        view.copy_(t_cpu)
        with torch.cuda.device(device):
            out = view.to(device=device, non_blocking=True)
            st.ev = torch.cuda.Event()
            st.ev.record()
    return out if dtype is None or out.dtype == dtype else out.to(dtype)


def _refuse_empty(B):
    """Before any step counter moves: kge_train_step refuses B == 0 (the reference's batch iterator never yields an
    empty batch, data_utils.py:176-196), and a refused call must leave adam_t / lazy_t where they were."""
    if B <= 0:
        raise ValueError("empty batch")


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


class PaddedTable:
    """(rows, cols) fp32 table stored with a row stride that is a multiple of 4 floats
    (16-byte rows for 128-bit access); `view` is the logical tensor."""

    def __init__(self, value, device):
        value = torch.as_tensor(np.asarray(value) if not torch.is_tensor(value) else value,
                                dtype=torch.float32)
        self.shape = tuple(value.shape)
        rows = self.shape[0]
        cols = int(np.prod(self.shape[1:]))
        self.rows, self.cols, self.ld = rows, cols, (cols + 3) // 4 * 4
        self.storage = torch.zeros((rows, self.ld), dtype=torch.float32, device=device)
        self.storage[:, :cols] = value.reshape(rows, cols).to(device)
        self.m = None
        self.v = None
        self.wn = None            # shadow: normalise(storage) maintained by the training step (kge_tables.wn_valid)
        self.wn_valid = False
        self.last = None          # row-lazy replay: Adam step of every stored row (kge_tables.lazy_t)
        self.owner = None         # engine to flush before the table is read from outside the library

    def touch(self):
        """Call after writing `storage` / `view` from outside the library: the shadow is stale."""
        self.wn_valid = False

    @property
    def view(self):
        if self.owner is not None and getattr(self.owner, "_w_stale", False):
            raise RuntimeError("data-parallel run with dp_w_private: call Engine.dp_sync_weights() on every rank before "
                               "reading the weights")
        if self.last is not None and self.owner is not None:
            self.owner.flush()
        return self.storage[:, :self.cols].reshape(self.shape)

    def ensure_slots(self, acc_init=0.0):
        if self.m is None:
            self.m = torch.zeros_like(self.storage)
            self.v = torch.zeros_like(self.storage)
            if acc_init:
                self.v[:, :self.cols] = acc_init      # padding stays zero


class StepGraph:
    """kge_step_graph of an Engine: one captured training step replayed per batch.  `launch` takes the batch from
    host (pinned) or device memory; with `next_triples` the following batch is staged and grouped under this step,
    and the launch that trains on it (same `next_step`) passes no data at all."""

    def __init__(self, eng, B, seed, lookahead=True, dp=None):
        if eng.lazy:
            raise RuntimeError("step graphs cover dense tables (not row-lazy)")
        self.eng, self.B, self.seed, self.dp = eng, int(B), int(seed), dp
        dt = torch.int64 if eng.cfg.id_dtype == 1 else torch.int32
        self.stage = [torch.empty((B, 3), dtype=dt, device=eng.device) for _ in range(2 if lookahead else 1)]
        self.loss = torch.zeros(1, dtype=torch.float32, device=eng.device)
        self.ws = torch.empty(eng._workspace(B, 0).numel(), dtype=torch.uint8, device=eng.device)   # the graph owns its workspace
        self.T = eng._tables(with_opt=True)
        self.h = C.c_void_p()
        nxt = _ptr(self.stage[1]) if lookahead else C.c_void_p(0)
        if dp is None:
            L.check(L.load().kge_step_graph_create(eng.ctx, C.byref(eng.cfg), C.byref(self.T), C.byref(eng.sampler),
                                                   _ptr(self.stage[0]), nxt, self.B, self.seed, _ptr(self.loss),
                                                   _ptr(self.ws), self.ws.numel(), C.byref(self.h)))
        else:
            from . import dp as _dp
            rank, world_size = _dp.world() if getattr(eng, "_dp_world", None) is None else eng._dp_world
            dpd = L.Dp(world=world_size, rank=rank, B_global=dp[0], i_base=dp[1])
            push = (L.Grads * 2)(eng._peer.push(0), eng._peer.push(1))
            slots = (L.Grads * 2)(eng._peer.slots(0), eng._peer.slots(1))
            L.check(L.load().kge_step_graph_create_dp(eng.ctx, C.byref(eng.cfg), C.byref(self.T), C.byref(eng.sampler),
                                                      C.byref(dpd), push, slots, C.byref(eng._peer.peer_flags()),
                                                      _ptr(self.stage[0]), nxt, self.B, self.seed, _ptr(self.loss),
                                                      _ptr(self.ws), self.ws.numel(), C.byref(self.h)))

    def launch(self, triples, step, next_triples=None, next_step=None, host_loss=None):
        """triples / next_triples: contiguous (B, 3) id tensors on the host (pinned for an asynchronous copy) or
        on the device; host_loss: pinned float tensor the loss is copied to (valid once the stream is synchronised).
        Returns the device loss tensor."""
        eng = self.eng
        eng.adam_t += 1
        T = eng._tables(with_opt=True)
        eng._lazy_dirty = False
        a_tr = C.c_void_p(triples.data_ptr()) if triples is not None else C.c_void_p(0)
        a_nx = C.c_void_p(next_triples.data_ptr()) if next_triples is not None else C.c_void_p(0)
        a_hl = C.c_void_p(host_loss.data_ptr()) if host_loss is not None else C.c_void_p(0)
        ns = step + 1 if next_step is None else next_step
        if self.dp is None:
            L.check(L.load().kge_step_graph_launch(self.h, C.byref(T), a_tr, step, eng.adam_t, a_nx, ns, a_hl, _stream()))
        else:                       # the engine owns the exchange's step parity and barrier epoch (shared with train_step_dp)
            par = eng._dp_steps & 1
            eng._dp_steps += 1
            eng._dp_epoch = getattr(eng, "_dp_epoch", 0) + 2
            L.check(L.load().kge_step_graph_launch_dp(self.h, C.byref(T), a_tr, step, eng.adam_t, eng._dp_epoch - 1, par,
                                                      a_nx, ns, a_hl, _stream()))
            eng._w_stale = bool(getattr(eng._peer, "w_private", False))
        eng._sync_shadow(T)
        return self.loss

    def wait_loss(self):
        """Host waits until the last launch's loss is in its `host_loss` buffer (kge_step_graph_wait_loss): with a
        pinned buffer on one GPU that is before the step's update kernels have finished."""
        L.check(L.load().kge_step_graph_wait_loss(self.h))

    def phases(self):
        """-> ({phase: ms}, total ms) of the last launch (graph created under set_option(OPT_GRAPH_PROFILE, 1))."""
        ms = (C.c_float * L.N_PHASES)()
        tot = C.c_float()
        L.check(L.load().kge_step_graph_phases(self.h, ms, C.byref(tot)))
        return {L.PHASES[i]: float(ms[i]) for i in range(L.N_PHASES)}, float(tot.value)

    def __del__(self):
        try:
            if self.h:
                L.load().kge_step_graph_destroy(self.h)
                self.h = None
        except Exception:
            pass


class Engine:
    def __init__(self, cfg_kwargs, tables, device="cuda:0", lazy=False):
        """cfg_kwargs: fields of kge_cfg; tables: {slot: array-like}.  lazy: row-lazy replay of the
        dense per-batch passes on the entity tables (kge_abi.h, kge_tables.lazy_t) — the training step
        then touches the looked-up rows only; results are bit-identical to the dense step."""
        self.device = torch.device(device)
        self.ctx = _ctx(self.device.index or 0)
        self.cfg = L.Cfg(abi_version=L.KGE_ABI_VERSION, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-7,
                         clip_norm=5.0, id_dtype=1, score_p=2.0, rotate_scale=1.0)
        self.acc_init = float(cfg_kwargs.get("acc_init", 0.0))      # Adagrad: initial_accumulator_value of the v slot
        for k, v in cfg_kwargs.items():
            if k != "acc_init":
                setattr(self.cfg, k, v)
        self.tab = {s: PaddedTable(v, self.device) for s, v in tables.items()}
        for pt in self.tab.values():
            pt.owner = self
        self.lazy = bool(lazy)
        self.lazy_t = 0
        self._lazy_dirty = False
        if self.lazy:
            for s in (L.T_ENT, L.T_ENT_PROJ):
                if s in self.tab:
                    self.tab[s].last = torch.zeros(self.tab[s].rows, dtype=torch.int32, device=self.device)
                    self.tab[s].owner = self
        elif self.cfg.constraint and self.cfg.model in (L.UM, L.TRANSE, L.DISTMULT):
            self.tab[L.T_ENT].wn = torch.zeros_like(self.tab[L.T_ENT].storage)
        self.sampler = L.Sampler()
        self._sampler_keep = []
        self._ws = None
        self.adam_t = 0

    # ---- struct packing -----------------------------------------------------------
    def _tables(self, with_opt):
        """kge_tables of the current storages, rebuilt only when a pointer changed (tables are
        updated in place; Adam slots appear once)."""
        if with_opt:
            for pt in self.tab.values():
                pt.ensure_slots(self.acc_init)
        key = (bool(with_opt),) + tuple((pt.storage.data_ptr(), 0 if pt.m is None else pt.m.data_ptr())
                                        for pt in self.tab.values())
        cache = getattr(self, "_tables_cache", None)
        if cache is None:
            cache = self._tables_cache = {}
        if key in cache:
            T = cache[key]
            for s, pt in self.tab.items():
                T.wn_valid[s] = int(pt.wn_valid)
            T.lazy_t = self.lazy_t
            return T
        T = L.Tables()
        for s, pt in self.tab.items():
            T.t[s] = L.Table(w=pt.storage.data_ptr(), m=pt.m.data_ptr() if pt.m is not None else None,
                             v=pt.v.data_ptr() if pt.v is not None else None,
                             wn=pt.wn.data_ptr() if pt.wn is not None else None,
                             last=pt.last.data_ptr() if pt.last is not None else None,
                             rows=pt.rows, cols=pt.cols, ld=pt.ld)
            T.wn_valid[s] = int(pt.wn_valid)
        T.lazy_t = self.lazy_t
        if len(cache) > 8:
            cache.clear()
        cache[key] = T
        return T

    def _sync_shadow(self, T):
        """The library updates kge_tables.wn_valid; mirror it on the tables."""
        for s, pt in self.tab.items():
            pt.wn_valid = bool(T.wn_valid[s])
        self.lazy_t = int(T.lazy_t)

    def flush(self):
        """Row-lazy tables: replay every row to the current step (kge_flush_rows); no-op otherwise."""
        if getattr(self, "_w_stale", False):
            raise RuntimeError("data-parallel run with dp_w_private: call Engine.dp_sync_weights() on every rank first")
        if not self.lazy or not self._lazy_dirty:
            return
        self._lazy_dirty = False
        T = self._tables(with_opt=True)
        L.check(L.load().kge_flush_rows(self.ctx, C.byref(self.cfg), C.byref(T), _stream()))

    def set_uniform_pool(self, pool=None):
        self.cfg.sampler = L.NS_UNIFORM
        self.sampler = L.Sampler()
        if pool is not None:
            p = torch.as_tensor(np.asarray(pool), dtype=torch.int64).to(self.device)
            self._sampler_keep = [p]
            self.sampler.pool = p.data_ptr()
            self.sampler.pool_len = p.numel()

    def set_typed_pools(self, ent_type, type_off, type_members, ent_pos):
        self.cfg.sampler = L.NS_TYPED
        ts = [torch.as_tensor(np.asarray(a), dtype=torch.int64).to(self.device)
              for a in (ent_type, type_off, type_members, ent_pos)]
        self._sampler_keep = ts
        self.sampler = L.Sampler(ent_type=ts[0].data_ptr(), type_off=ts[1].data_ptr(),
                                 type_members=ts[2].data_ptr(), ent_pos=ts[3].data_ptr(),
                                 n_types=ts[1].numel() - 1)

    def _workspace(self, B, Q):
        memo = getattr(self, "_ws_memo", None)
        if memo is None:
            memo = self._ws_memo = {}
        if (B, Q) in memo and self._ws is not None and self._ws.numel() >= memo[(B, Q)]:
            return self._ws
        need = C.c_size_t()
        L.check(L.load().kge_workspace_bytes(C.byref(self.cfg), B, Q, C.byref(need)))
        memo[(B, Q)] = need.value
        if self._ws is None or self._ws.numel() < need.value:
            self._ws = torch.empty(need.value, dtype=torch.uint8, device=self.device)
        return self._ws

    def k_eff(self):
        k = self.cfg.negative_ratio
        return 2 * (k // 2) if self.cfg.corrupt_side == L.SIDE_HT else k

    def _ids(self, x):
        dt = torch.int64 if self.cfg.id_dtype == 1 else torch.int32
        if torch.is_tensor(x) and x.dtype == dt and x.device == self.device and x.is_contiguous():
            return x
        if torch.is_tensor(x):
            return x.to(device=self.device, dtype=dt).contiguous()
        return torch.as_tensor(np.ascontiguousarray(x), dtype=dt).to(self.device)

    # ---- calls ----------------------------------------------------------------------
    def sample_negatives(self, triples, seed, step):
        X = self._ids(triples)
        out = torch.empty(X.shape[0] * self.k_eff(), dtype=torch.int64, device=self.device)
        L.check(L.load().kge_sample_negatives(self.ctx, C.byref(self.cfg), C.byref(self.sampler), _ptr(X),
                                              X.shape[0], seed, step, _ptr(out), _stream()))
        return out

    def train_step(self, triples, seed, step, is_train=True, neg_override=None, loss_out=None,
                   want_negatives=False, next_triples=None, next_step=None):
        """One batch through kge_train_step.  next_triples (a device tensor of the engine's id dtype that stays
        unchanged until the call that trains on it) / next_step: the following batch, whose gradient records are
        grouped while this step runs (kge_train_step_next); the results do not depend on it."""
        X = self._ids(triples)
        _refuse_empty(X.shape[0])
        if not is_train:
            self.flush()                    # (also refuses stale data-parallel replicas)
        if next_triples is not None and is_train and neg_override is None and not want_negatives and not self.lazy:
            B = X.shape[0]
            ws = self._workspace(B, 0)
            if loss_out is None:
                loss_out = torch.empty(1, dtype=torch.float32, device=self.device)
            Xn = self._ids(next_triples)
            if Xn.shape[0] != B:
                Xn = None                               # ragged last batch: grouped inline by its own call
            self.adam_t += 1
            T = self._tables(with_opt=True)
            L.check(L.load().kge_train_step_next(self.ctx, C.byref(self.cfg), C.byref(T), C.byref(self.sampler), _ptr(X), B,
                                                 seed, step, self.adam_t, _ptr(loss_out), _ptr(Xn),
                                                 step + 1 if next_step is None else next_step, _ptr(ws), ws.numel(),
                                                 _stream()))
            self._sync_shadow(T)
            self._keep_next = Xn                        # the library reads it when the next step is enqueued
            return loss_out
        B = X.shape[0]
        ws = self._workspace(B, 0)
        if loss_out is None:
            loss_out = torch.empty(1, dtype=torch.float32, device=self.device)
        neg = None
        if neg_override is not None:
            neg = torch.as_tensor(neg_override, dtype=torch.int64).to(self.device).contiguous()
        neg_out = torch.empty(B * self.k_eff(), dtype=torch.int64, device=self.device) if want_negatives else None
        if is_train:
            self.adam_t += 1
        T = self._tables(with_opt=bool(is_train) or self.lazy)
        self._lazy_dirty = bool(is_train)
        L.check(L.load().kge_train_step(self.ctx, C.byref(self.cfg), C.byref(T), C.byref(self.sampler), _ptr(X), B,
                                        seed, step, max(self.adam_t, 1), _ptr(neg), int(bool(is_train)),
                                        _ptr(loss_out), _ptr(neg_out), _ptr(ws), ws.numel(), _stream()))
        self._sync_shadow(T)
        return (loss_out, neg_out) if want_negatives else loss_out

    def train_step_host(self, host_triples, seed, step, host_loss, is_train=True, next_host=None, next_step=None):
        """Batch and loss in (pinned) host memory: H2D copy, step and D2H of the loss are enqueued by
        ONE library call (kge_train_step_host); the caller synchronises the stream to read the loss.
        next_host: the following batch (pinned host tensor, same shape): copied and grouped under this step
        (kge_train_step_host_next); the call that trains on it copies nothing."""
        B = host_triples.shape[0]
        _refuse_empty(B)
        if next_host is not None and is_train and not self.lazy and next_host.shape[0] == B:
            dt = torch.int64 if self.cfg.id_dtype == 1 else torch.int32
            if host_triples.dtype != dt or next_host.dtype != dt or not host_triples.is_contiguous() or not next_host.is_contiguous():
                raise ValueError(f"host batches must be contiguous {dt} tensors")
            st2 = getattr(self, "_stage2", None)
            if st2 is None or st2[0].shape[0] != B:
                st2 = self._stage2 = [torch.empty((B, 3), dtype=dt, device=self.device) for _ in range(2)]
                self._stage2_cur = 0
                self._stage_loss = torch.zeros(1, dtype=torch.float32, device=self.device)
            cur = self._stage2_cur
            ws = self._workspace(B, 0)
            self.adam_t += 1
            T = self._tables(with_opt=True)
            L.check(L.load().kge_train_step_host_next(self.ctx, C.byref(self.cfg), C.byref(T), C.byref(self.sampler),
                                                      C.c_void_p(host_triples.data_ptr()), _ptr(st2[cur]), B, seed, step,
                                                      self.adam_t, _ptr(self._stage_loss), C.c_void_p(host_loss.data_ptr()),
                                                      C.c_void_p(next_host.data_ptr()), _ptr(st2[1 - cur]),
                                                      step + 1 if next_step is None else next_step, _ptr(ws), ws.numel(),
                                                      _stream()))
            self._sync_shadow(T)
            self._stage2_cur = 1 - cur
            return host_loss
        dt = torch.int64 if self.cfg.id_dtype == 1 else torch.int32
        if host_triples.dtype != dt or not host_triples.is_contiguous():
            raise ValueError(f"host_triples must be a contiguous {dt} tensor")
        st = getattr(self, "_stage", None)
        if st is None or st.shape[0] < B:
            st = self._stage = torch.empty((B, 3), dtype=dt, device=self.device)
            self._stage_loss = torch.zeros(1, dtype=torch.float32, device=self.device)
        ws = self._workspace(B, 0)
        if is_train:
            self.adam_t += 1
        T = self._tables(with_opt=bool(is_train) or self.lazy)
        self._lazy_dirty = bool(is_train)
        L.check(L.load().kge_train_step_host(self.ctx, C.byref(self.cfg), C.byref(T), C.byref(self.sampler),
                                             C.c_void_p(host_triples.data_ptr()), _ptr(st), B, seed, step,
                                             max(self.adam_t, 1), int(bool(is_train)), _ptr(self._stage_loss),
                                             C.c_void_p(host_loss.data_ptr()), _ptr(ws), ws.numel(), _stream()))
        self._sync_shadow(T)
        return host_loss

    def step_graph(self, B, seed, lookahead=True):
        """A training step for batches of B triples captured as a CUDA graph (kge_step_graph): see StepGraph."""
        return StepGraph(self, B, seed, lookahead)

    # ---- data-parallel step: grad -> one all-reduce -> apply ---------------
    def grad_buffer(self):
        from . import dp as _dp
        return _dp.GradBuffer({s: (pt.rows, pt.ld) for s, pt in self.tab.items()}, self.device)

    def _grads_struct(self, gb):
        G = L.Grads()
        for s in self.tab:
            G.dense[s] = gb.view(s).data_ptr()
        return G

    def train_grad(self, gb, triples_local, seed, step, B_global, i_base, rank=0, world_size=1,
                   neg_override=None, want_negatives=False):
        """kge_train_grad: this rank's summed gradients and [norm^2 x4, loss] into `gb`."""
        X = self._ids(triples_local)
        B = X.shape[0]
        ws = self._workspace(B, 0)
        dpd = L.Dp(world=world_size, rank=rank, B_global=B_global, i_base=i_base)
        neg = None
        if neg_override is not None:
            neg = torch.as_tensor(neg_override, dtype=torch.int64).to(self.device).contiguous()
        neg_out = torch.empty(B * self.k_eff(), dtype=torch.int64, device=self.device) if want_negatives else None
        T = self._tables(with_opt=False)
        if isinstance(gb, L.Grads):            # peer staging (dp.PeerStaging.push): scalars travel with it
            G, scal = gb, None
        else:
            G, scal = self._grads_struct(gb), gb.scal
        L.check(L.load().kge_train_grad(self.ctx, C.byref(self.cfg), C.byref(T), C.byref(self.sampler),
                                        C.byref(dpd), _ptr(X), B, seed, step, _ptr(neg), C.byref(G), _ptr(scal),
                                        _ptr(neg_out), _ptr(ws), ws.numel(), _stream()))
        return neg_out

    def train_apply(self, gb, B, loss_out=None):
        """kge_train_apply with the (rank-summed) buffer `gb`; B = the local batch the workspace was sized for."""
        ws = self._workspace(B, 0)
        if loss_out is None:
            loss_out = torch.empty(1, dtype=torch.float32, device=self.device)
        self.adam_t += 1
        T = self._tables(with_opt=True)
        if isinstance(gb, L.Grads):            # peer staging (dp.PeerStaging.slots)
            G, scal = gb, None
        else:
            G, scal = self._grads_struct(gb), gb.scal
        L.check(L.load().kge_train_apply(self.ctx, C.byref(self.cfg), C.byref(T), C.byref(G), _ptr(scal), B,
                                         self.adam_t, _ptr(loss_out), _ptr(ws), ws.numel(), _stream()))
        self._sync_shadow(T)
        return loss_out

    def _dp_setup(self, group=None):
        """First data-parallel call: pick the exchange and set up its peer memory (collective)."""
        from . import dp as _dp
        rank, world_size = _dp.world()
        if getattr(self, "_dp_kind", None) is None:
            dense_reg = bool(self.cfg.constraint) and self.cfg.model in (L.TRANSH, L.RESCAL)
            self._dp_kind, self._peer, self._gbuf, self._dp_steps = _dp.exchange_kind(world_size, dense_reg), None, None, 0
            if self.cfg.model in (L.SE, L.TRANSR, L.RESCAL):
                # matrix models: the relation matrices' gradients are dense [n_rel x d d'] tables, exchanged by ONE
                # all-reduce of the flat buffer (kge_train_grad / kge_train_apply through grads.dense)
                self._dp_kind = "nccl"
            if self._dp_kind in ("peer", "owner"):
                layout = {s: (pt.rows, pt.ld) for s, pt in self.tab.items()}
                wn_slots = None
                if self._dp_kind == "owner":
                    wn_slots = [s for s, pt in self.tab.items() if pt.wn is not None]
                self._peer = _dp.PeerStaging.symmetric(layout, self.device, group, wn_slots=wn_slots,
                                                       w_private=getattr(self, "dp_w_private", False))
                if self._peer is None:
                    self._dp_kind = "nccl"
                elif self._dp_kind == "owner":
                    self.adopt_peer_tables(self._peer.ptabs)
                    self._peer.barrier()
            if self._dp_kind == "nccl":
                self._gbuf = self.grad_buffer()
        return rank, world_size

    def dp_sync_weights(self, group=None):
        """Data-parallel runs with dp_w_private = True (set before the first data-parallel step): the new w rows
        stay with their owners and only wn travels per step; this gathers w from the owners (collective: every rank
        must call it) — before evaluation, validation batches, checkpoints or any read of the weights."""
        peer = getattr(self, "_peer", None)
        if peer is not None and getattr(self, "_w_stale", False):
            torch.cuda.current_stream().synchronize()
            peer.sync_weights(self.tab, group)
            self._w_stale = False

    def step_graph_dp(self, B, seed, B_global, i_base, group=None, lookahead=True):
        """The data-parallel step (kge_train_step_dp_next) captured as a CUDA graph; None when the exchange is not
        carried by peer memory (NCCL all-reduce path: the collective is not capturable here)."""
        self._dp_setup(group)
        if self._dp_kind not in ("peer", "owner"):
            return None
        return StepGraph(self, B, seed, lookahead, dp=(B_global, i_base))

    def train_step_dp(self, triples_local, seed, step, B_global, i_base, loss_out=None, neg_override=None, group=None,
                      next_triples=None, next_step=None):
        """One data-parallel step; every rank of the process group must call it.  next_triples: this rank's
        share of the following step's batch (look-ahead grouping, kge_train_step_dp_next)."""
        rank, world_size = self._dp_setup(group)
        if self._dp_kind in ("peer", "owner") and neg_override is None and os.environ.get("KGE_DP_ONE_CALL", "1") != "0":
            # grad -> barrier kernel -> apply -> barrier kernel enqueued by ONE library call (kge_train_step_dp)
            par = self._dp_steps & 1
            self._dp_steps += 1
            X = self._ids(triples_local)
            B = X.shape[0]
            ws = self._workspace(B, 0)
            if loss_out is None:
                loss_out = torch.empty(1, dtype=torch.float32, device=self.device)
            dpd = L.Dp(world=world_size, rank=rank, B_global=B_global, i_base=i_base)
            self.adam_t += 1
            T = self._tables(with_opt=True)
            self._dp_epoch = getattr(self, "_dp_epoch", 0) + 2
            Xn = self._ids(next_triples) if next_triples is not None else None
            if Xn is not None and Xn.shape[0] != B:
                Xn = None
            L.check(L.load().kge_train_step_dp_next(self.ctx, C.byref(self.cfg), C.byref(T), C.byref(self.sampler), C.byref(dpd),
                                                    _ptr(X), B, seed, step, self.adam_t, C.byref(self._peer.push(par)),
                                                    C.byref(self._peer.slots(par)), C.byref(self._peer.peer_flags()),
                                                    self._dp_epoch - 1, _ptr(loss_out), _ptr(Xn),
                                                    step + 1 if next_step is None else next_step, _ptr(ws), ws.numel(),
                                                    _stream()))
            self._sync_shadow(T)
            self._keep_next = Xn
            self._w_stale = bool(getattr(self._peer, "w_private", False))
            return loss_out
        if self._dp_kind == "peer":
            # gradient kernels store into every rank's staging slot; one barrier; sum of the slots
            par = self._dp_steps & 1
            self._dp_steps += 1
            self.train_grad(self._peer.push(par), triples_local, seed, step, B_global, i_base, rank, world_size, neg_override)
            self._peer.barrier()
            return self.train_apply(self._peer.slots(par), len(triples_local), loss_out)
        if self._dp_kind == "owner":
            # row sums to the rows' owners; barrier; every owner updates its rows and stores them into
            # every replica; barrier (the next step gathers rows other ranks wrote)
            par = self._dp_steps & 1
            self._dp_steps += 1
            self.train_grad(self._peer.push(par), triples_local, seed, step, B_global, i_base, rank, world_size, neg_override)
            self._peer.barrier()
            out = self.train_apply(self._peer.slots(par), len(triples_local), loss_out)
            self._peer.barrier()
            return out
        self.train_grad(self._gbuf, triples_local, seed, step, B_global, i_base, rank, world_size, neg_override)
        self._gbuf.allreduce(group)                       # the step's single collective
        return self.train_apply(self._gbuf, len(triples_local), loss_out)

    def check_peers_alive(self):
        """Data-parallel runs over peer memory: raise if a cross-GPU barrier timed out (syncs the device)."""
        peer = getattr(self, "_peer", None)
        if peer is not None:
            peer.check_alive()

    def adopt_peer_tables(self, ptabs):
        """Move the weight tables (and shadow tables) into peer-mapped memory (dp.PeerTables)."""
        for s, pt in self.tab.items():
            region = ptabs.view("w", s)
            region.copy_(pt.storage)
            pt.storage = region
            if pt.wn is not None and ("wn", s) in ptabs.off:
                rn = ptabs.view("wn", s)
                rn.copy_(pt.wn)
                pt.wn = rn
        self._tables_cache = {}
        self._ptabs = ptabs

    def score(self, h, r, t):
        self.flush()
        h, r, t = self._ids(h), self._ids(r), self._ids(t)
        out = torch.empty(h.numel(), dtype=torch.float32, device=self.device)
        T = self._tables(False)
        L.check(L.load().kge_score(self.ctx, C.byref(self.cfg), C.byref(T), _ptr(h), _ptr(r), _ptr(t),
                                   h.numel(), _ptr(out), _stream()))
        return out

    def constrain_table(self, slot, op):
        self.flush()
        T = self._tables(False)
        L.check(L.load().kge_constrain_table(self.ctx, C.byref(T), slot, op, _stream()))
        self.tab[slot].touch()

    @staticmethod
    def _filter_struct(f_off, f_ids):
        """kge_filter from (f_off, f_ids): f_off int64[Q + 1] = a packed CSR, or int64[2, Q] = begin | end
        ranges into f_ids (what filter.FilterIndex.lookup returns)."""
        flt = L.Filter()
        if f_off is not None:
            if f_off.dim() == 2:
                flt.off, flt.end = f_off[0].data_ptr(), f_off[1].data_ptr()
            else:
                flt.off = f_off.data_ptr()
            flt.ids = f_ids.data_ptr()
        return flt

    def filter_index(self, positive_X, side):
        """Device-side known-positive index of positive_X for one corrupt side (kge_filter_index_build)."""
        from . import filter as _filter
        return _filter.FilterIndex(self.ctx, self.cfg, positive_X, side, self.device)

    def rank_counts(self, queries, side, f_off=None, f_ids=None, ent_lo=0, ent_hi=None):
        self.flush()
        Qx = self._ids(queries)
        Q = Qx.shape[0]
        cnt = torch.zeros(Q, dtype=torch.int32, device=self.device)
        ws = self._workspace(0, Q)
        flt = self._filter_struct(f_off, f_ids)
        T = self._tables(False)
        L.check(L.load().kge_rank(self.ctx, C.byref(self.cfg), C.byref(T), _ptr(Qx), Q, side, C.byref(flt),
                                  ent_lo, self.cfg.n_ent if ent_hi is None else ent_hi, _ptr(cnt), _ptr(ws),
                                  ws.numel(), _stream()))
        return cnt

    def rank_counts_static(self, queries, side, f_off=None, f_ids=None, ent_lo=0, ent_hi=None, out=None):
        """kge_rank_lp_static: rank_counts without a workspace — the scratch is the library's static device
        array (include/kge_abi.h); `out` (zeroed int32[Q]) lets the caller keep the call allocation-free."""
        self.flush()
        Qx = self._ids(queries)
        Q = Qx.shape[0]
        cnt = torch.zeros(Q, dtype=torch.int32, device=self.device) if out is None else out
        flt = self._filter_struct(f_off, f_ids)
        T = self._tables(False)
        L.check(L.load().kge_rank_lp_static(self.ctx, C.byref(self.cfg), C.byref(T), _ptr(Qx), Q, side, C.byref(flt),
                                            ent_lo, self.cfg.n_ent if ent_hi is None else ent_hi, _ptr(cnt), _stream()))
        return cnt

    def set_option(self, option, value):
        """Implementation switch of the context (shared by every engine on the device)."""
        L.check(L.load().kge_set_option(self.ctx, option, int(value)))

    def profile_begin(self, max_calls):
        L.check(L.load().kge_profile_begin(self.ctx, max_calls))

    def profile_end(self):
        """-> ({phase: summed ms}, n_calls)"""
        ms = (C.c_float * L.N_PHASES)()
        n = C.c_int()
        L.check(L.load().kge_profile_end(self.ctx, ms, C.byref(n)))
        return {L.PHASES[i]: float(ms[i]) for i in range(L.N_PHASES)}, n.value
