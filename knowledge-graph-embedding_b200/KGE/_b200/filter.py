"""Known-positive filter index for ranking (kge_filter CSR).

Replaces the two full NumPy scans of positive_X per query at BaseModel.py:646-650 by
one sort of positive_X and two binary searches per query (torch ops on the GPU: this
is index plumbing outside the scoring kernels)."""
import torch


def build_filter(positive_X, queries, side, n_ent, device):
    """side: 'h' (candidates replace the head; match on (r, t)) or 't' (match on (h, r)).
    Returns (off[Q+1], ids[total]) int64 device tensors, ids[off[q]:off[q+1]] = entities to
    exclude for query q (the query's own true entity included when it is in positive_X)."""
    P = torch.as_tensor(positive_X).to(device=device, dtype=torch.int64)
    P = torch.unique(P, dim=0)          # duplicate triples would be subtracted twice by the CSR-subtraction kernels
    Qx = torch.as_tensor(queries).to(device=device, dtype=torch.int64)
    fs, cs = (2, 0) if side == "h" else (0, 2)
    pkey = P[:, 1] * n_ent + P[:, fs]
    order = torch.argsort(pkey, stable=True)
    pkey, pent = pkey[order].contiguous(), P[order, cs].contiguous()
    qkey = (Qx[:, 1] * n_ent + Qx[:, fs]).contiguous()
    lo = torch.searchsorted(pkey, qkey, right=False)
    hi = torch.searchsorted(pkey, qkey, right=True)
    lens = hi - lo
    off = torch.zeros(Qx.shape[0] + 1, dtype=torch.int64, device=device)
    off[1:] = torch.cumsum(lens, 0)
    total = int(off[-1].item())
    if total == 0:
        return off, torch.zeros(1, dtype=torch.int64, device=device)
    src = torch.arange(total, device=device) - torch.repeat_interleave(off[:-1], lens) \
        + torch.repeat_interleave(lo, lens)
    return off, pent[src].contiguous()
