"""Known-positive filter index for ranking (kge_filter), built by the library on the device.

Replaces the two full NumPy scans of positive_X PER QUERY at BaseModel.py:646-650 by ONE sort of
positive_X per (positive_X, side) (`kge_filter_index_build`: composite 64-bit keys -> radix sort ->
unique -> split) and two binary searches per query (`kge_filter_lookup`).  The ranking kernels read
the index's own sorted entity array through per-query [begin, end) ranges: nothing is gathered, and
no step synchronises with the host.  torch only owns the buffers."""
import ctypes as C

import numpy as np
import torch

from . import lib as L


def _side(side):
    if side in ("h", L.SIDE_H):
        return L.SIDE_H
    if side in ("t", L.SIDE_T):
        return L.SIDE_T
    raise ValueError("side must be 'h' or 't'")


class FilterIndex:
    """Index of `positive_X` for one corrupt side.  `ctx`, `cfg`: the engine's context and kge_cfg
    (n_ent, n_rel and id_dtype are what matters).  Build once per evaluate() call; `lookup` per block."""

    def __init__(self, ctx, cfg, positive_X, side, device):
        self.ctx, self.cfg, self.side, self.device = ctx, cfg, _side(side), torch.device(device)
        dt = torch.int64 if cfg.id_dtype == 1 else torch.int32
        P = positive_X if torch.is_tensor(positive_X) else torch.as_tensor(np.ascontiguousarray(np.asarray(positive_X)))
        from . import engine as _engine
        P = _engine.upload(P, self.device, dt).reshape(-1, 3).contiguous()
        self.n_pos = int(P.shape[0])
        if self.n_pos == 0:
            raise ValueError("positive_X is empty: pass positive_X=None for unfiltered ranks")
        ib, wb = C.c_size_t(), C.c_size_t()
        L.check(L.load().kge_filter_index_bytes(C.byref(cfg), self.n_pos, C.byref(ib), C.byref(wb)))
        self.index = torch.empty(ib.value, dtype=torch.uint8, device=self.device)
        ws = torch.empty(wb.value, dtype=torch.uint8, device=self.device)
        st = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        L.check(L.load().kge_filter_index_build(ctx, C.byref(cfg), C.c_void_p(P.data_ptr()), self.n_pos, self.side,
                                                C.c_void_p(self.index.data_ptr()), ib.value,
                                                C.c_void_p(ws.data_ptr()), wb.value, st))
        ws.record_stream(torch.cuda.current_stream(self.device))      # freed only after the build has run
        P.record_stream(torch.cuda.current_stream(self.device))
        off_key = 256
        off_ent = off_key + (8 * self.n_pos + 255) // 256 * 256
        self.ents = self.index[off_ent:off_ent + 8 * self.n_pos].view(torch.int64)     # sorted candidate entities

    def lookup(self, queries):
        """-> (ranges int64[2, Q] = begin | end, ents): ids to exclude for query q are ents[begin[q]:end[q]]
        (the query's own true entity included when its triple is in positive_X).  `queries`: device tensor
        (Q, 3) of the cfg's id dtype."""
        Q = int(queries.shape[0])
        be = torch.empty((2, Q), dtype=torch.int64, device=self.device)
        st = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        L.check(L.load().kge_filter_lookup(self.ctx, C.byref(self.cfg), C.c_void_p(self.index.data_ptr()), self.n_pos,
                                           C.c_void_p(queries.data_ptr()), Q, self.side, C.c_void_p(be[0].data_ptr()),
                                           C.c_void_p(be[1].data_ptr()), None, st))
        return be, self.ents

    def n_unique(self):
        """Number of distinct in-range triples in the index (reads the device header: synchronises)."""
        return int(self.index[32:40].view(torch.int64).item())


def build_filter(positive_X, queries, side, n_ent, device, n_rel=None):
    """One-shot convenience (tests, scripts): index + lookup.  Returns (ranges[2, Q], ents) as
    Engine.rank_counts takes them.  `n_rel` defaults to 1 + the largest relation id seen."""
    from . import engine as _engine
    device = torch.device(device)
    P = np.asarray(positive_X.cpu() if torch.is_tensor(positive_X) else positive_X).reshape(-1, 3)
    Qx = np.asarray(queries.cpu() if torch.is_tensor(queries) else queries).reshape(-1, 3)
    if n_rel is None:
        n_rel = int(max(P[:, 1].max(), Qx[:, 1].max())) + 1
    cfg = L.Cfg(abi_version=L.KGE_ABI_VERSION, model=L.TRANSE, d_ent=1, d_rel=1, id_dtype=1, score_p=1.0,
                n_ent=int(n_ent), n_rel=int(n_rel))
    idx = FilterIndex(_engine._ctx(device.index or 0), cfg, P.astype(np.int64), side, device)
    qd = torch.as_tensor(np.ascontiguousarray(Qx.astype(np.int64))).to(device)
    return idx.lookup(qd)
