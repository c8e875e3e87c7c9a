"""Where KGEModel.train()'s per-step time goes on the C2 shape: the loop rebuilt from its parts (device batch stream,
step graph with look-ahead, per-step loss copy), host enqueue time vs. time to completion for each combination."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "knowledge-graph-embedding_b200"), os.path.join(ROOT, "tests")]
from KGE._b200 import engine as E, spec
from KGE.data_utils import DeviceBatchStream
from oracle import ref_ops as O, ref_models as RM

dev = torch.device("cuda", 0)
N, R, d, B, k, n_train = 14951, 1345, 200, 1024, 64, 483142
rng = np.random.default_rng(0)
train = np.stack([rng.integers(0, N, n_train), rng.integers(0, R, n_train), rng.integers(0, N, n_train)], 1).astype(np.int64)
ref = RM.MODELS["DistMult"](N, R, d, None, score_fn=None, loss_fn=O.BinaryCrossEntropyLoss(), constraint=True, constraint_weight=1.0)
W = ref.init_weights(5)
cfg = spec.make_cfg("DistMult", N, R, d, None, score_fn=None, loss_fn=O.BinaryCrossEntropyLoss(), negative_ratio=k, corrupt_side="h+t",
                    constraint=True, constraint_weight=1.0, rotate_limit=ref.limit)
eng = E.Engine(cfg, {spec.SLOTS["DistMult"][kk]: v for kk, v in W.data.items()}, device=dev)
graph = eng.step_graph(B, 1)
n_steps = 472
stream = DeviceBatchStream(train, B, shuffle=True, seed=1, device="cuda:0")
fixed = [next(stream) for _ in range(8)]
t_loss = torch.zeros(n_steps, dtype=torch.float32, device=dev)
step = [0]


def run(name, use_stream, copy_loss, reps=4):
    res = []
    for _ in range(reps):
        nxt = next(stream) if use_stream else fixed[0]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for b in range(n_steps):
            if use_stream:
                cur, nxt = nxt, next(stream)
            else:
                cur, nxt = fixed[b & 7], fixed[(b + 1) & 7]
            out = graph.launch(cur, step[0], next_triples=nxt, next_step=step[0] + 1)
            if copy_loss:
                t_loss[b:b + 1].copy_(out)
            step[0] += 1
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        res.append(((t1 - t0) / n_steps * 1e6, (t2 - t0) / n_steps * 1e6))
    print("%-44s host enqueue %6.1f us/step, to completion %6.1f us/step" % (name, min(r[0] for r in res), min(r[1] for r in res)))


run("graph.launch only, fixed batches", False, False)
run("+ per-step loss copy", False, True)
run("+ DeviceBatchStream", True, False)
run("+ both (the training loop)", True, True)
