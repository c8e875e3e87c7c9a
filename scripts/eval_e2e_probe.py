"""Where the end-to-end filtered evaluation of bench.py spends its time (C2 shape): device time per segment
(CUDA events) and host time of the whole call."""
import os, sys, time, json, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "knowledge-graph-embedding_b200"), os.path.join(ROOT, "tests")]
from KGE._b200 import engine as E, spec
from oracle import ref_ops as O, ref_models as RM
dev = torch.device("cuda", 0)
N, R, d, Q = 14951, 1345, 200, 98304
rng = np.random.default_rng(7)
ref = RM.MODELS["DistMult"](N, R, d, None, score_fn=None, loss_fn=O.BinaryCrossEntropyLoss(), constraint=True, constraint_weight=1.0)
W = ref.init_weights(5)
cfg = spec.make_cfg("DistMult", N, R, d, None, score_fn=None, loss_fn=O.BinaryCrossEntropyLoss(), negative_ratio=64, corrupt_side="h+t",
                    constraint=True, constraint_weight=1.0, rotate_limit=ref.limit)
eng = E.Engine(cfg, {spec.SLOTS["DistMult"][k]: v for k, v in W.data.items()}, device=dev)
train = np.stack([rng.integers(0, N, 483142), rng.integers(0, R, 483142), rng.integers(0, N, 483142)], 1)
q_all = train[rng.integers(0, len(train), Q)]
pos_host = torch.from_numpy(train).pin_memory(); q_host = torch.from_numpy(q_all).pin_memory()
out_host = torch.zeros(Q, dtype=torch.int32).pin_memory()
def ev(): return torch.cuda.Event(enable_timing=True)
acc = {}
for rep in range(6):
    for s, sid in (("h", 0), ("t", 1)):
        e = [ev() for _ in range(6)]
        t0 = time.perf_counter()
        e[0].record(); pos_dev = pos_host.to(dev, non_blocking=True); q_dev = q_host.to(dev, non_blocking=True)
        e[1].record(); idx = eng.filter_index(pos_dev, s)
        e[2].record(); f = idx.lookup(q_dev)
        e[3].record(); c = eng.rank_counts(q_dev, sid, f[0], f[1])
        e[4].record(); out_host.copy_(c, non_blocking=True)
        e[5].record()
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        if rep >= 2:
            for n, (a, b) in zip(("h2d", "index_build", "lookup", "rank", "d2h"), zip(e[:-1], e[1:])):
                acc[n] = acc.get(n, 0) + a.elapsed_time(b)
            acc["host_enqueue_ms"] = acc.get("host_enqueue_ms", 0) + (t1 - t0) * 1e3
            acc["wall_ms"] = acc.get("wall_ms", 0) + (t2 - t0) * 1e3
print(json.dumps({k: round(v / 8, 3) for k, v in acc.items()}))
# bench.py's loop: both sides back to back, one synchronisation per repetition
outs = {s: torch.zeros(Q, dtype=torch.int32).pin_memory() for s in ("h", "t")}
for rep in range(5):
    t0 = time.perf_counter()
    marks = []
    for s, sid in (("h", 0), ("t", 1)):
        pos_dev = pos_host.to(dev, non_blocking=True); q_dev = q_host.to(dev, non_blocking=True)
        marks.append(time.perf_counter())
        idx = eng.filter_index(pos_dev, s)
        marks.append(time.perf_counter())
        f = idx.lookup(q_dev)
        c = eng.rank_counts(q_dev, sid, f[0], f[1])
        marks.append(time.perf_counter())
        outs[s].copy_(c, non_blocking=True)
        marks.append(time.perf_counter())
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    print("rep", rep, "wall_ms", round((t1 - t0) * 1e3, 2), "host marks ms", [round((m - t0) * 1e3, 2) for m in marks])

# one evaluate()-shaped call (bench.py's end-to-end leg): 59 071 queries per side in blocks of 16 384, one index per call
Qe, blk = 59071, 16384
q_e = train[rng.integers(0, len(train), Qe)]
qe_host = torch.from_numpy(q_e).pin_memory()
oute = torch.zeros(Qe, dtype=torch.int32).pin_memory()
for rep in range(5):
    for s, sid in (("h", 0), ("t", 1)):
        names, evs = [], []
        def mark(n):
            e = ev(); e.record(); names.append(n); evs.append(e)
        t0 = time.perf_counter()
        mark("start")
        pos_dev = pos_host.to(dev, non_blocking=True); q_dev = qe_host.to(dev, non_blocking=True)
        mark("h2d")
        idx = eng.filter_index(pos_dev, s)
        mark("index")
        for lo in range(0, Qe, blk):
            qb = q_dev[lo:lo + blk]
            f = idx.lookup(qb)
            mark("lookup")
            c = eng.rank_counts(qb, sid, f[0], f[1])
            mark("rank")
            oute[lo:lo + blk].copy_(c, non_blocking=True)
            mark("d2h")
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        if rep == 4:
            seg = {}
            for n, a, b in zip(names[1:], evs[:-1], evs[1:]):
                seg[n] = round(seg.get(n, 0) + a.elapsed_time(b), 3)
            print("evaluate() call side", s, "device segments ms", seg, "sum", round(evs[0].elapsed_time(evs[-1]), 3),
                  "host enqueue", round((t1 - t0) * 1e3, 3), "wall", round((t2 - t0) * 1e3, 3))
