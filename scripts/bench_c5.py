"""C5: synthetic N = 5 M entities x R = 1000 relations, TransE d = 128, L1, hinge,
uniform 'h+t' k = 32 — the one configuration whose tables (2.56 GB + Adam slots) do not sit in L2, so
the training step is a true HBM gather / scatter.  One GPU: row-lazy step (kge_tables.lazy_t) vs the
dense step the reference's semantics imply, both constraint settings; filtered ranking of Q queries
against all 5 M entities.  One JSON line per measurement.
    python scripts/bench_c5.py [--n-ent 5000000] [--B 8192 65536] [--steps 30] [--dense-steps 4] [--eval-q 4096]"""
import argparse, json, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "knowledge-graph-embedding_b200"), os.path.join(ROOT, "tests")]
from KGE._b200 import engine as E, spec
from KGE._b200.filter import build_filter
from oracle import ref_ops as O

ap = argparse.ArgumentParser()
ap.add_argument("--n-ent", type=int, default=5_000_000)
ap.add_argument("--n-rel", type=int, default=1000)
ap.add_argument("--d", type=int, default=128)
ap.add_argument("--B", type=int, nargs="+", default=[8192, 65536])
ap.add_argument("--steps", type=int, default=30)
ap.add_argument("--dense-steps", type=int, default=4)
ap.add_argument("--eval-q", type=int, default=4096)
ap.add_argument("--profile-only", action="store_true", help="lazy constraint=True B[0] only, few steps (for ncu)")
a = ap.parse_args()
dev = torch.device("cuda", 0)
N, R, d, k = a.n_ent, a.n_rel, a.d, 32
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0}
rng = np.random.default_rng(12345)
wrel = 1.0 / np.arange(1, R + 1)
wrel /= wrel.sum()


def triples(n):
    return np.stack([rng.integers(0, N, n), rng.choice(R, size=n, p=wrel), rng.integers(0, N, n)], 1).astype(np.int64)


def engine(constraint, lazy):
    g = torch.Generator().manual_seed(12345)
    lim = 6.0 / np.sqrt(d)
    W = {0: (torch.rand((N, d), generator=g) * 2 - 1) * lim, 2: (torch.rand((R, d), generator=g) * 2 - 1) * lim}
    cfg = spec.make_cfg("TransE", N, R, d, score_fn=O.LpDistance(1), loss_fn=O.PairwiseHingeLoss(1.0), negative_ratio=k,
                        corrupt_side="h+t", constraint=constraint)
    return E.Engine(cfg, W, device=dev, lazy=lazy)


def time_steps(eng, B, steps, warm):
    batches = [torch.from_numpy(triples(B)).to(dev) for _ in range(8)]
    loss = torch.zeros(1, device=dev)
    for i in range(warm):
        eng.train_step(batches[i % 8], 3, i, loss_out=loss)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    torch.cuda.synchronize()
    for i in range(steps):
        ev[i][0].record(); eng.train_step(batches[(warm + i) % 8], 3, warm + i, loss_out=loss); ev[i][1].record()
    torch.cuda.synchronize()
    ms = sorted(x.elapsed_time(y) for x, y in ev)
    return sum(ms) / len(ms), ms[len(ms) // 2], float(loss.item())


def report(name, B, ms, med, loss, extra=None):
    alg = B * (7 * 4 * ((2 + k) * d + d) + 24)          # 125 452 B per positive at d = 128, k' = 32
    out = {"config": name, "B": B, "ms_per_step": round(ms, 4), "ms_median": round(med, 4),
           "train_triples_per_s": round(B * (1 + k) / (ms * 1e-3)), "algorithmic_GBps": round(alg / (ms * 1e-3) / 1e9, 1),
           "hbm_frac": round(alg / (ms * 1e-3) / 1e9 / peaks["hbm_gbs"], 3), "final_loss": loss}
    out.update(extra or {})
    print(json.dumps(out), flush=True)


if a.profile_only:
    eng = engine(True, True)
    ms, med, loss = time_steps(eng, a.B[0], 4, 3)
    report("C5 lazy constraint=True (profile run)", a.B[0], ms, med, loss)
    sys.exit(0)

for constraint in (False, True):
    for lazy in (True, False):
        eng = engine(constraint, lazy)
        for B in a.B:
            steps = a.steps if lazy else a.dense_steps
            ms, med, loss = time_steps(eng, B, steps, 3)
            ph = None
            eng.profile_begin(8)
            time_steps(eng, B, 6, 0)
            phases, n = eng.profile_end()
            ph = {kk: round(v / max(n, 1), 4) for kk, v in phases.items() if v > 0}
            report(f"C5 TransE d={d} L1 hinge h+t k=32 N={N} constraint={constraint} {'row-lazy' if lazy else 'dense'}",
                   B, ms, med, loss, {"phase_ms": ph})
        if lazy:
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record(); eng.flush(); a1.record(); torch.cuda.synchronize()
            print(json.dumps({"config": f"kge_flush_rows N={N} constraint={constraint}", "ms": round(a0.elapsed_time(a1), 3)}), flush=True)
        if lazy and constraint:
            train = triples(200_000)
            Xq = train[rng.integers(0, len(train), a.eval_q)]
            qd = torch.from_numpy(Xq).to(dev)
            filt = {s: build_filter(train, Xq, s, N, dev) for s in ("h", "t")}
            for s in ("h", "t"):
                eng.rank_counts(qd[:256], 0 if s == "h" else 1)
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(); a0.record()
            for s in ("h", "t"):
                eng.rank_counts(qd, 0 if s == "h" else 1, *filt[s])
            a1.record(); torch.cuda.synchronize()
            sec = a0.elapsed_time(a1) * 1e-3
            print(json.dumps({"config": f"C5 filtered eval N={N} d={d} L1 (tile kernel)", "queries": 2 * a.eval_q,
                              "eval_queries_per_s": round(2 * a.eval_q / sec), "ms": round(sec * 1e3, 2),
                              "lp_ops_per_s_T": round(2 * a.eval_q * N * d * 3 / sec / 1e12, 2)}), flush=True)
        del eng
        torch.cuda.empty_cache()
