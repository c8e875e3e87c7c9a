"""C1-shaped filtered evaluation (TransE d=50, LpDistance(1), N=14505, R=237, both sides):
tiled Lp kernel vs the generic warp-per-candidate kernel.  Prints one JSON line."""
import json, os, sys, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "knowledge-graph-embedding_b200"), os.path.join(ROOT, "tests")]
from KGE._b200 import engine as E, spec
from KGE._b200.filter import build_filter
from oracle import ref_ops as O

N, R, d, Q = 14505, 237, 50, 17526
rng = np.random.default_rng(3)
g = torch.Generator().manual_seed(1)
lim = 6 / d ** 0.5
W = {"ent_emb": (torch.rand((N, d), generator=g) * 2 - 1) * lim, "rel_emb": (torch.rand((R, d), generator=g) * 2 - 1) * lim}
cfg = spec.make_cfg("TransE", N, R, d, score_fn=O.LpDistance(1), loss_fn=O.PairwiseHingeLoss(1.0), negative_ratio=1,
                    corrupt_side="t", constraint=True)
eng = E.Engine(cfg, {spec.SLOTS["TransE"][k]: v for k, v in W.items()})
train = np.stack([rng.integers(0, N, 272115), rng.integers(0, R, 272115), rng.integers(0, N, 272115)], 1)
Xq = train[rng.integers(0, len(train), Q)]
filt = {s: build_filter(train, Xq, s, N, eng.device) for s in ("h", "t")}
qd = torch.from_numpy(Xq).to(eng.device)

def run(reps):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    out = {}
    for s in ("h", "t"):
        out[s] = eng.rank_counts(qd, 0 if s == "h" else 1, *filt[s])
    torch.cuda.synchronize(); a.record()
    for _ in range(reps):
        for s in ("h", "t"):
            eng.rank_counts(qd, 0 if s == "h" else 1, *filt[s])
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps, out

ms_tile, c_tile = run(5)
os.environ["KGE_RANK_GENERIC"] = "1"
ms_gen, c_gen = run(1)
os.environ.pop("KGE_RANK_GENERIC")
diff = max(int((c_tile[s].long() - c_gen[s].long()).abs().max()) for s in ("h", "t"))
ops = 2.0 * 2 * Q * N * d                    # sub + abs-add per (query, entity, dim), both sides
print(json.dumps({"workload": "C1 eval: TransE d=50 L1, N=14505, 17526 triples x 2 sides, filter = 272115 train triples",
                  "tile_ms": ms_tile, "tile_queries_per_s": 2 * Q / (ms_tile * 1e-3),
                  "generic_ms": ms_gen, "generic_queries_per_s": 2 * Q / (ms_gen * 1e-3),
                  "speedup": ms_gen / ms_tile, "max_count_diff_vs_generic": diff,
                  "alu_ops_per_s": ops / (ms_tile * 1e-3),
                  "fp32_alu_peak_ops_per_s": 148 * 128 * 1.965e9,
                  "frac_of_fp32_issue_peak": ops / (ms_tile * 1e-3) / (148 * 128 * 1.965e9)}))
