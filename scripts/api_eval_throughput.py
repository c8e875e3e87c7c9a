"""Throughput of KGEModel.evaluate() itself on the C2 shape: 59 071 queries per call (the FB15k test split's size),
positive_X = the 483 142 training triples, numpy arrays in, metric dict out — wall clock per call."""
import os, sys, time, tempfile
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "knowledge-graph-embedding_b200")]
from KGE import loss
from KGE.models.semantic_based.DistMult import DistMult

N, R, n_train, B, k, d, Q = 14951, 1345, 483142, 1024, 64, 200, 59071
rng = np.random.default_rng(0)
train = np.stack([rng.integers(0, N, n_train), rng.integers(0, R, n_train), rng.integers(0, N, n_train)], 1).astype(np.int64)
test = train[rng.integers(0, n_train, Q)]
meta = {"ind2ent": [str(i) for i in range(N)], "ind2rel": [str(i) for i in range(R)]}
m = DistMult(embedding_params={"embedding_size": d}, negative_ratio=k, corrupt_side="h+t", loss_fn=loss.BinaryCrossEntropyLoss())
with tempfile.TemporaryDirectory() as tmp:
    m.train(train_X=train, val_X=None, metadata=meta, epochs=1, batch_size=B, seed=1, log_path=tmp)
for side in ("h", "t"):
    m.evaluate(test, side, positive_X=train)
torch.cuda.synchronize()
ts = []
for rep in range(5):
    t0 = time.perf_counter()
    for side in ("h", "t"):
        out = m.evaluate(test, side, positive_X=train)
    ts.append(time.perf_counter() - t0)
tr = []
for rep in range(5):
    t0 = time.perf_counter()
    for side in ("h", "t"):
        m._rank_block(test, side, train)
    tr.append(time.perf_counter() - t0)
print("of which the ranks (upload, index, kge_rank blocks, read back): %.2f ms" % (min(tr) * 1e3))
best = min(ts)
print("evaluate() both sides, ms per pair of calls:", [round(x * 1e3, 2) for x in ts])
print("best: %.2f ms = %.2f M queries/s through the public API; mrr %.4f" % (best * 1e3, 2 * Q / best / 1e6, out["mean_reciprocal_rank"]))
