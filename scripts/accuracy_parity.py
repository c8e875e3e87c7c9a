"""Accuracy parity on the bundled FB15k-237 (BASELINE.json configs[0]; north_star: "matching MRR/Hits@10"):
TransE L1 d = 50, pairwise hinge, uniform 1:1 negatives, Adam — trained for the same seeded batch stream and the
same Philox negatives ONCE on the GPU path (libkge_b200 through the step graph / plain calls) and ONCE on the CPU
oracle (oracle/ref_step.run_single_batch: the reference's math, TF optimizer semantics), then filtered MRR / Hits@k
of the same validation queries, each side ranking with its own code (kge_rank vs the oracle's get_rank).

    python scripts/accuracy_parity.py [--epochs 3] [--queries 1000] [--out gpurun_out/accuracy_parity.json]

The two trajectories differ only by floating-point summation order (1e-7 per step), so losses agree to ~1e-5 and
the metrics to the third decimal; the ranks of ONE set of weights are compared exactly elsewhere
(tests/test_gpu_fullsize.py)."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "knowledge-graph-embedding_b200"), os.path.join(ROOT, "tests")]


def run(epochs=3, n_queries=1000, B=4096, seed=2024, threads=None, log=print):
    from helpers import make_pair, tables_of
    from oracle import ref_ops as O, ref_step as RS
    from KGE._b200.filter import build_filter
    g = np.load(os.path.join(ROOT, "tests", "golden", "data_bundled.npz"))
    train, valid = g["fb15k237_train"].astype(np.int64), g["fb15k237_valid"].astype(np.int64)
    N, R, d, k, side = 14505, 237, 50, 1, "t"
    if threads:
        torch.set_num_threads(threads)
    ref, W, eng = make_pair("TransE", N, R, d, score=O.LpDistance(1), loss=O.PairwiseHingeLoss(1.0), k=k, side=side, seed=5)
    opt = RS.TFAdam()
    n_b = len(train) // B                                    # whole batches only (both sides see the same stream)
    rng = np.random.default_rng(seed)
    G = eng.step_graph(B, seed)
    loss_gpu, loss_cpu, t_gpu, t_cpu, step = [], [], 0.0, 0.0, 0
    for ep in range(epochs):
        perm = rng.permutation(len(train))
        batches = [train[perm[b * B:(b + 1) * B]] for b in range(n_b)]
        dev = [torch.from_numpy(x).cuda() for x in batches]
        lg = torch.zeros(n_b, device="cuda")
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for b in range(n_b):
            lg[b] = G.launch(dev[b], step + b, next_triples=dev[b + 1] if b + 1 < n_b else None)[0]
        torch.cuda.synchronize()
        t_gpu += time.perf_counter() - t0
        t0 = time.perf_counter()
        lc = []
        for b in range(n_b):
            ids, ke = RS.sample_negatives(seed, step + b, batches[b], k, side, N)
            l, _ = RS.run_single_batch(ref, W, opt, batches[b], ids, ke, side, True)
            lc.append(l)
        t_cpu += time.perf_counter() - t0
        step += n_b
        loss_gpu.append(float(lg.mean().item()))
        loss_cpu.append(float(np.mean(lc)))
        log("epoch %d: mean batch loss gpu %.6f cpu %.6f" % (ep, loss_gpu[-1], loss_cpu[-1]))
    # filtered ranks of the same validation queries, positive_X = train U valid (BaseModel.py:646-654)
    qr = np.random.default_rng(1)
    Q = valid[qr.choice(len(valid), n_queries, replace=False)]
    posX = np.concatenate([train, valid])
    out = {}
    Wg = tables_of(eng, "TransE")
    dW = {kk: float((Wg[kk] - W.data[kk]).abs().max()) for kk in Wg}
    ranks = {"gpu": [], "cpu": []}
    t0 = time.perf_counter()
    for sid, sname in ((0, "h"), (1, "t")):
        off, ids = build_filter(posX, Q, sname, N, eng.device)
        ranks["gpu"].append(eng.rank_counts(Q, sid, off, ids).cpu().numpy().astype(np.float64) + 1)
    torch.cuda.synchronize()
    t_rank_gpu = time.perf_counter() - t0
    t0 = time.perf_counter()
    for sname in ("h", "t"):
        ranks["cpu"].append(np.array([RS.get_rank(ref, W, x, posX, sname) for x in Q], dtype=np.float64))
    t_rank_cpu = time.perf_counter() - t0
    for kk in ranks:
        r = np.concatenate(ranks[kk])
        out[kk] = {"mrr": float(np.mean(1 / r)), "mean_rank": float(r.mean()), "median_rank": float(np.median(r)),
                   "hits1": float(np.mean(r <= 1)), "hits3": float(np.mean(r <= 3)), "hits10": float(np.mean(r <= 10))}
    rg, rc = np.concatenate(ranks["gpu"]), np.concatenate(ranks["cpu"])
    res = {"dataset": "bundled fb15k_237 (272 115 train / 17 526 valid triples, 14 505 entities, 237 relations)",
           "model": "TransE L1 d=50, PairwiseHingeLoss(1), uniform negative_ratio=1 corrupt_side=t, Adam(tf defaults), B=4096",
           "epochs": epochs, "steps": step, "queries": 2 * n_queries, "filter": "train U valid",
           "gpu": out["gpu"], "cpu_oracle": out["cpu"],
           "abs_diff": {m: abs(out["gpu"][m] - out["cpu"][m]) for m in out["gpu"]},
           "ranks_identical_fraction": float(np.mean(rg == rc)), "rank_abs_diff_max": float(np.abs(rg - rc).max()),
           "epoch_loss_gpu": loss_gpu, "epoch_loss_cpu": loss_cpu,
           "epoch_loss_max_rel_diff": float(max(abs(a - b) / max(abs(b), 1e-12) for a, b in zip(loss_gpu, loss_cpu))),
           "max_abs_weight_diff": dW,
           "train_seconds": {"gpu": t_gpu, "cpu_oracle": t_cpu, "cpu_threads": torch.get_num_threads()},
           "rank_seconds": {"gpu": t_rank_gpu, "cpu_oracle": t_rank_cpu}}
    return res


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=3)
    ap.add_argument("--queries", type=int, default=1000)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    r = run(a.epochs, a.queries)
    s = json.dumps(r, indent=1)
    print(s)
    if a.out:
        os.makedirs(os.path.dirname(a.out), exist_ok=True)
        open(a.out, "w").write(s)
