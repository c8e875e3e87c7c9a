"""Per-phase device times (library CUDA events) of one configuration:
    python scripts/phases_cfg.py TransR|SE|RESCAL|TransD|RotatE"""
import json, os, sys, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "knowledge-graph-embedding_b200"), os.path.join(ROOT, "tests")]
from KGE._b200 import engine as E, spec
from oracle import ref_ops as O, ref_models as RM
model = sys.argv[1] if len(sys.argv) > 1 else "TransR"
N, R, B = 14505, 237, 1024
if model == "RotatE":      # C3
    d, k, score, loss = 500, 256, O.LpDistance(1), O.SelfAdversarialNegativeSamplingLoss(3.0, 1.0)
else:                      # C4 shape
    d, k, score, loss = 100, 10, O.LpDistancePow(2), O.PairwiseHingeLoss(1.0)
dev = torch.device("cuda", 0)
rng = np.random.default_rng(7)
ref = RM.MODELS[model](N, R, d, d, score_fn=score, loss_fn=loss, constraint=True, constraint_weight=1.0)
W = ref.init_weights(5)
if model == "TransR":
    W.data["rel_proj"] = W.data["rel_proj"] + (torch.rand(W.data["rel_proj"].shape) - 0.5) * 0.2
cfg = spec.make_cfg(model, N, R, d, d, score_fn=score, loss_fn=loss, negative_ratio=k, corrupt_side="h+t", constraint=True,
                    rotate_limit=ref.limit)
eng = E.Engine(cfg, {spec.SLOTS[model][kk]: v for kk, v in W.data.items()}, device=dev)
train = np.stack([rng.integers(0, N, 272115), rng.integers(0, R, 272115), rng.integers(0, N, 272115)], 1)
batches = [torch.from_numpy(train[rng.integers(0, len(train), B)]).to(dev) for _ in range(16)]
loss_dev = torch.zeros(1, device=dev)
for i in range(5):
    eng.train_step(batches[i % 16], 3, i, loss_out=loss_dev)
eng.profile_begin(40)
for i in range(40):
    eng.train_step(batches[i % 16], 3, 10 + i, loss_out=loss_dev)
torch.cuda.synchronize()
ph, n = eng.profile_end()
print(json.dumps({"model": model, "us_per_step": {kk: round(v / n * 1e3, 1) for kk, v in ph.items() if v > 0}}))
