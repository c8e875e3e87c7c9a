"""Training-step and filtered-eval throughput of the SURVEY.md 8(d) configurations C1..C4 on one
GPU (synthetic triples of each config's shape; L2 flushed between timed steps).  One JSON line per
configuration.  bench.py stays the headline (C2); this fills BASELINE.md section 4."""
import json, os, sys, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "knowledge-graph-embedding_b200"), os.path.join(ROOT, "tests")]
from KGE._b200 import engine as E, spec
from KGE._b200.filter import build_filter
from oracle import ref_ops as O, ref_models as RM

dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def typed_pools(eng, N, n_types, rng):
    t = rng.integers(0, n_types, N)
    order = np.argsort(t, kind="stable")
    off = np.concatenate([[0], np.cumsum(np.bincount(t, minlength=n_types))])
    pos = np.empty(N, dtype=np.int64); pos[order] = np.arange(N) - off[t[order]]
    eng.set_typed_pools(t, off, order, pos)


def run(name, model, N, R, d, dr, score, loss, k, side, B, typed=0, steps=60, evalQ=4096, constraint=True):
    rng = np.random.default_rng(7)
    ref = RM.MODELS[model](N, R, d, dr, score_fn=score, loss_fn=loss, constraint=constraint, constraint_weight=1.0)
    W = ref.init_weights(5)
    if model == "TransR":
        W.data["rel_proj"] = W.data["rel_proj"] + (torch.rand(W.data["rel_proj"].shape) - 0.5) * 0.2
    cfg = spec.make_cfg(model, N, R, d, dr, score_fn=score, loss_fn=loss, negative_ratio=k, corrupt_side=side,
                        constraint=constraint, constraint_weight=1.0, rotate_limit=ref.limit)
    eng = E.Engine(cfg, {spec.SLOTS[model][kk]: v for kk, v in W.data.items()}, device=dev)
    if typed:
        typed_pools(eng, N, typed, rng)
    train = np.stack([rng.integers(0, N, 272115), rng.integers(0, R, 272115), rng.integers(0, N, 272115)], 1)
    batches = [torch.from_numpy(train[rng.integers(0, len(train), B)]).to(dev) for _ in range(16)]
    loss_dev = torch.zeros(1, device=dev)
    for i in range(5):
        eng.train_step(batches[i % 16], 3, i, loss_out=loss_dev)

    def timed(fn):
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        torch.cuda.synchronize()
        for i in range(steps):
            flush.zero_(); ev[i][0].record(); fn(i); ev[i][1].record()
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in ev) / steps
    ms_plain = timed(lambda i: eng.train_step(batches[i % 16], 3, 10 + i, loss_out=loss_dev))
    # the product's loop: captured step graph, next batch grouped under the step
    G = eng.step_graph(B, 3)
    for i in range(4):
        G.launch(batches[i % 16], 500 + i, next_triples=batches[(i + 1) % 16])
    ms = timed(lambda i: G.launch(batches[i % 16], 1000 + i, next_triples=batches[(i + 1) % 16]))
    loss_dev = G.loss
    k_eff = eng.k_eff()
    Xq = train[rng.integers(0, len(train), evalQ)]
    filt = {s: build_filter(train, Xq, s, N, dev) for s in ("h", "t")}
    qd = torch.from_numpy(Xq).to(dev)
    for s in ("h", "t"):
        eng.rank_counts(qd, 0 if s == "h" else 1, *filt[s])
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for s in ("h", "t"):
        eng.rank_counts(qd, 0 if s == "h" else 1, *filt[s])
    b.record(); torch.cuda.synchronize()
    print(json.dumps({"config": name, "B": B, "k_eff": k_eff, "train_us_per_step": round(ms * 1e3, 1), "train_us_per_step_plain_calls": round(ms_plain * 1e3, 1),
                      "train_triples_per_s": round(B * (1 + k_eff) / (ms * 1e-3)),
                      "eval_queries_per_s": round(2 * evalQ / (a.elapsed_time(b) * 1e-3)),
                      "final_loss": float(loss_dev.item())}), flush=True)


hinge = O.PairwiseHingeLoss(1.0)
run("C1 TransE d=50 L1 hinge 't' k=1 B=4096", "TransE", 14505, 237, 50, None, O.LpDistance(1), hinge, 1, "t", 4096, evalQ=17526)
run("C1 TransE d=50 L1 hinge 't' k=1 B=65536", "TransE", 14505, 237, 50, None, O.LpDistance(1), hinge, 1, "t", 65536, evalQ=17526)
run("C2 DistMult d=200 BCE 'h+t' k=64 B=1024", "DistMult", 14951, 1345, 200, None, None, O.BinaryCrossEntropyLoss(), 64, "h+t", 1024, evalQ=16384)
run("C3 RotatE d=500 L1 self-adv 'h+t' k=256 B=1024", "RotatE", 14505, 237, 500, None, O.LpDistance(1),
    O.SelfAdversarialNegativeSamplingLoss(3.0, 1.0), 256, "h+t", 1024, steps=20, evalQ=2048)
run("C4 TransR d=100 L2^2 hinge typed(3) 'h+t' k=10 B=1024", "TransR", 14505, 237, 100, 100, O.LpDistancePow(2), hinge, 10, "h+t", 1024, typed=3, evalQ=16384)
run("C4 TransD d=100 L2^2 hinge typed(32) 'h+t' k=10 B=1024", "TransD", 14505, 237, 100, 100, O.LpDistancePow(2), hinge, 10, "h+t", 1024, typed=32, evalQ=16384)
# C5 shape at one GPU's share: 1/5 of the entities (1 M x 128, w + m + v = 1.5 GB), B = 65536 / 8
# (C5 itself is bench.py's c5 leg: row-lazy / sharded tables)
if False: run("C5-lite TransE d=128 L1 hinge 'h+t' k=32 B=8192 N=1M constraint=False", "TransE", 1000000, 1000, 128, None, O.LpDistance(1), hinge,
    32, "h+t", 8192, steps=20, evalQ=1024, constraint=False)
