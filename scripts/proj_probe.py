"""One filtered ranking pass of a C4-shaped model (TransR / TransD / TransH / SE) at the product's block size
(16 384 queries): wall time per side, and — under `ncu --metrics gpu__time_duration.sum` — the launch list of
the projected-table path (csrc/rank_proj_tile.cuh)."""
import os, sys, json, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "knowledge-graph-embedding_b200"), os.path.join(ROOT, "tests")]
from KGE._b200 import engine as E, spec
from KGE._b200.filter import build_filter
from oracle import ref_ops as O, ref_models as RM

model = sys.argv[1] if len(sys.argv) > 1 else "TransR"
Q = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
N, R, d = 14505, 237, 100
dev = torch.device("cuda", 0)
rng = np.random.default_rng(7)
ref = RM.MODELS[model](N, R, d, d, score_fn=O.LpDistancePow(2), loss_fn=O.PairwiseHingeLoss(1.0), constraint=True, constraint_weight=1.0)
W = ref.init_weights(5)
cfg = spec.make_cfg(model, N, R, d, d, score_fn=O.LpDistancePow(2), loss_fn=O.PairwiseHingeLoss(1.0), negative_ratio=10,
                    corrupt_side="h+t", constraint=True, constraint_weight=1.0, rotate_limit=ref.limit)
eng = E.Engine(cfg, {spec.SLOTS[model][k]: v for k, v in W.data.items()}, device=dev)
train = np.stack([rng.integers(0, N, 272115), rng.integers(0, R, 272115), rng.integers(0, N, 272115)], 1)
Xq = train[rng.integers(0, len(train), Q)]
qd = torch.from_numpy(Xq).to(dev)
out = {"model": model, "Q": Q}
for s in ("h", "t"):
    f = build_filter(train, Xq, s, N, dev)
    eng.rank_counts(qd, 0 if s == "h" else 1, *f)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    eng.rank_counts(qd, 0 if s == "h" else 1, *f)
    b.record(); torch.cuda.synchronize()
    out["ms_" + s] = round(a.elapsed_time(b), 3)
out["queries_per_s"] = round(2 * Q / ((out["ms_h"] + out["ms_t"]) * 1e-3))
print(json.dumps(out))
