"""Where the end-to-end step (host batch in, loss out, host synchronised every step) loses time against the
resident-input step: the step graph launched with device / host batches, with / without the loss copy and the sync."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "knowledge-graph-embedding_b200"), os.path.join(ROOT, "tests")]
from KGE._b200 import engine as E, spec
from oracle import ref_ops as O, ref_models as RM

dev = torch.device("cuda", 0)
N, R, d, B, k = 14951, 1345, 200, 1024, 64
rng = np.random.default_rng(0)
ref = RM.MODELS["DistMult"](N, R, d, None, score_fn=None, loss_fn=O.BinaryCrossEntropyLoss(), constraint=True, constraint_weight=1.0)
W = ref.init_weights(5)
cfg = spec.make_cfg("DistMult", N, R, d, None, score_fn=None, loss_fn=O.BinaryCrossEntropyLoss(), negative_ratio=k, corrupt_side="h+t",
                    constraint=True, constraint_weight=1.0, rotate_limit=ref.limit)
eng = E.Engine(cfg, {spec.SLOTS["DistMult"][kk]: v for kk, v in W.data.items()}, device=dev)
graph = eng.step_graph(B, 1)
nb = 16
host = [torch.from_numpy(np.stack([rng.integers(0, N, B), rng.integers(0, R, B), rng.integers(0, N, B)], 1).astype(np.int64)).pin_memory()
        for _ in range(nb)]
devb = [h.to(dev) for h in host]
loss_host = torch.zeros(1).pin_memory()
step = [0]
n_steps = 400


def run(name, batches, with_loss, sync):
    best = None
    for _ in range(4):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for b in range(n_steps):
            graph.launch(batches[b % nb], step[0], next_triples=batches[(b + 1) % nb], next_step=step[0] + 1,
                         host_loss=loss_host if with_loss else None)
            step[0] += 1
            if sync:
                torch.cuda.current_stream().synchronize()
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        r = ((t1 - t0) / n_steps * 1e6, (t2 - t0) / n_steps * 1e6)
        best = r if best is None or r[1] < best[1] else best
    print("%-58s host %6.1f us/step, to completion %6.1f us/step" % (name, best[0], best[1]))


run("device batches, no loss copy, no sync", devb, False, False)
run("host batches (H2D nodes in the graph), no loss copy, no sync", host, False, False)
run("host batches + loss D2H, no sync", host, True, False)
run("host batches + loss D2H + sync every step (bench e2e)", host, True, True)
run("device batches + loss D2H + sync every step", devb, True, True)
