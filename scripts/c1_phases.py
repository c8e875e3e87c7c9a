"""Graph timeline of the C1 step (TransE d=50 L1 hinge 't' k=1, B=4096) and of C2, from instrumented step graphs."""
import json, os, sys, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "knowledge-graph-embedding_b200"), os.path.join(ROOT, "tests")]
from KGE._b200 import engine as E, spec, lib as L
from oracle import ref_ops as O, ref_models as RM
dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def run(name, model, N, R, d, score, loss, k, side, B):
    rng = np.random.default_rng(7)
    ref = RM.MODELS[model](N, R, d, None, score_fn=score, loss_fn=loss, constraint=True, constraint_weight=1.0)
    W = ref.init_weights(5)
    cfg = spec.make_cfg(model, N, R, d, None, score_fn=score, loss_fn=loss, negative_ratio=k, corrupt_side=side, constraint=True, constraint_weight=1.0, rotate_limit=ref.limit)
    eng = E.Engine(cfg, {spec.SLOTS[model][kk]: v for kk, v in W.data.items()}, device=dev)
    batches = [torch.from_numpy(np.stack([rng.integers(0, N, B), rng.integers(0, R, B), rng.integers(0, N, B)], 1)).to(dev) for _ in range(16)]
    eng.set_option(L.OPT_GRAPH_PROFILE, 1)
    G = eng.step_graph(B, 3)
    eng.set_option(L.OPT_GRAPH_PROFILE, 0)
    acc, tot, n = {}, 0.0, 30
    for i in range(n + 4):
        flush.zero_()
        G.launch(batches[i % 16], i, next_triples=batches[(i + 1) % 16])
        ph, t = G.phases()
        if i >= 4:
            tot += t
            for kk, v in ph.items(): acc[kk] = acc.get(kk, 0) + v
    print(name, {kk: round(v / n * 1e3, 1) for kk, v in acc.items() if v > 0}, "total_us", round(tot / n * 1e3, 1), flush=True)
run("C1 B=4096", "TransE", 14505, 237, 50, O.LpDistance(1), O.PairwiseHingeLoss(1.0), 1, "t", 4096)
run("C1 B=65536", "TransE", 14505, 237, 50, O.LpDistance(1), O.PairwiseHingeLoss(1.0), 1, "t", 65536)
run("C2", "DistMult", 14951, 1345, 200, None, O.BinaryCrossEntropyLoss(), 64, "h+t", 1024)
