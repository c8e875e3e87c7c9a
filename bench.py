#!/usr/bin/env python
"""bench.py — train triples/s (pos+neg) and filtered-eval queries/s of the KGE hot path.

Workload = BASELINE.json configs[1] (C2): DistMult d=200, BinaryCrossEntropyLoss,
Uniform 'h+t' negative_ratio=64, batch 1024, FB15k shape (N=14951, R=1345), synthetic
Zipf triples (no network for the dataset), filtered ranking on both sides.

    python bench.py --gpus N --steps K --warmup W [--impl reference]

Prints ONE JSON line (see the task contract).  `value` = device-resident inputs,
`e2e` = host buffers through the C-ABI call with H2D/D2H inside the timed region,
`roofline` = the dominant kernel (k_step) timed live with CUDA events on its stream,
`cpu_baseline` = the CPU oracle (port of the reference math; TensorFlow is not
installable here) on a bounded sample of the same workload.
"""
import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "knowledge-graph-embedding_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

if os.environ.get("NCCL_DEBUG", "VERSION") == "VERSION":
    os.environ["NCCL_DEBUG"] = "WARN"          # keep stdout to the one JSON line (NCCL prints its version there)

import numpy as np
import torch

WORKLOAD = {"workload": "C2 DistMult d=200 BCE uniform h+t k=64 B=1024 FB15k-shape (N=14951,R=1345)",
            "model": "DistMult", "n_ent": 14951, "n_rel": 1345, "d": 200, "B": 1024, "k": 64,
            "side": "h+t", "n_train": 483142, "eval_Q": 16384, "eval_Q_e2e": 59071}
METRIC = "train triples/sec (pos+neg)"


def peaks():
    """HBM / bf16 peaks: the driver-written MEASURED_PEAKS.json.  TF32 dense and NVLink store bandwidth are not in
    that file: measured on the box by scripts/measure_peaks.py -> profiles/r02_extra_peaks.json (committed)."""
    out = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "src": "fallback"}
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        out = {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"], "src": "measured"}
    out["tf32_tflops"], out["tf32_src"] = out["bf16_tflops"] / 2, "tf32 dense assumed = bf16 cuBLAS burst / 2"
    x = os.path.join(ROOT, "profiles", "r02_extra_peaks.json")
    if os.path.exists(x):
        e = json.load(open(x))
        if e.get("tf32_tflops"):
            out["tf32_tflops"] = e["tf32_tflops"]
            out["tf32_src"] = "measured: cuBLAS TF32 8192^3 best of 10 (profiles/r02_extra_peaks.json)"
        out["nvlink_store_gbs"] = e.get("nvlink_store_gbs")
    return out


def roofline_all(share, fwd_bytes, upd_bytes, pk, world):
    """One entry per phase of the C2 step (library CUDA events, ms per step): the critical path is
    max(constraint + step, grouping chain) + update (+ apply and barriers when data parallel)."""
    W = WORKLOAD
    k_eff = 2 * (W["k"] // 2)
    n_rec = W["B"] * (2 + k_eff) + W["B"]
    n_keys = n_rec
    rec_bytes = 4 * W["d"] * n_rec
    rows = W["n_ent"] + W["n_rel"]
    items = [("k_step<MDistMult,2,true>", "step", fwd_bytes, "forward gather 4*((2+k')*d + d) B per positive + ids"),
             ("grouping chain: k_keys, k_scan, k_scatter, k_canon (side stream)", "sort", 28 * n_keys + 32 * rows,
              "12*n_keys (keys) + 16*n_keys (scatter + canon) + 32 B header per table row"),
             ("k_update_fast / k_update_heavy" + ("<3> (row sums to owners)" if world > 1 else ""), "update",
              (upd_bytes if world == 1 else 4 * W["d"] * rows) + rec_bytes,
              "6*4*rows*cols (w, m, v read + write) + the gradient records read back" if world == 1 else
              "gradient records read + one row sum written per table row"),
             ("k_dp_scales + k_update_fast<4> (owner apply, rows stored to every replica)", "apply",
              upd_bytes // max(world, 1) + 4 * W["d"] * rows, "6*4*rows/G*cols + (G-1)/G of a w (+ wn) table out over NVLink"),
             ("k_peer_barrier (first of two)", "barrier", 0, "latency only"),
             ("k_constrain (skipped when the shadow table is valid)", "constraint", 8 * W["d"] * W["n_ent"], "2*4*rows*cols")]
    out = []
    for name, ph, nbytes, what in items:
        ms = share.get(ph)
        if not ms:
            continue
        gbs = nbytes / (ms * 1e-3) / 1e9
        out.append({"kernel": name, "phase": ph, "avg_ms_per_step": ms, "bound": "hbm", "algorithmic_bytes": int(nbytes),
                    "achieved": gbs, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": gbs / pk["hbm_gbs"], "bytes_are": what})
    return out


def zipf_triples(n, N, R, seed):
    """h, t ~ Zipf(1.0) over a seeded permutation of entities, r ~ Zipf(1.0)."""
    rng = np.random.default_rng(seed)

    def zipf(count, size):
        w = 1.0 / np.arange(1, count + 1)
        return rng.permutation(count)[rng.choice(count, size=size, p=w / w.sum())]
    return np.stack([zipf(N, n), zipf(R, n), zipf(N, n)], 1).astype(np.int64)


class Clocks:
    """nvidia-smi sampler (B200_PROFILING.md recipe), running from before the warm-up to the end of the
    GPU-timed regions: every sample is taken while the bench keeps the GPU under its own load.  A
    reader thread collects the lines; `ensure()` keeps the workload running until a sample exists
    (nvidia-smi needs a second or more to start on an 8-GPU box, a 100-step region lasts 20 ms)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.idx, self.thread = [], None, gpu_index, None

    def start(self):
        import threading
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
        except Exception:
            self.proc = None
            return self

        def read():
            for line in self.proc.stdout:
                f = [x.strip() for x in line.split(",")]
                if len(f) >= 9:
                    self.rows.append(f)
        self.thread = threading.Thread(target=read, daemon=True)
        self.thread.start()
        return self

    def stop(self):
        if self.proc is None:
            return
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        if self.thread is not None:
            self.thread.join(timeout=2)
        self.proc = None

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for j, n in enumerate(names) if any(r[5 + j].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": float(self.rows[0][2]) if self.rows[0][2].replace(".", "").isdigit() else None,
                "samples": len(self.rows), "reasons": reasons}


def ncu_traffic(kernel_prefix):
    """dram bytes (read + write) per launch of a kernel from the committed ncu capture
    (profiles/r01_traffic.json), or None."""
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r01_traffic.json")) as fh:
            t = json.load(fh)["traffic_bytes_per_launch"]
        for name, v in t.items():
            if name.startswith(kernel_prefix):
                return v
    except Exception:
        pass
    return None


def init_tables(seed=12345):
    W = WORKLOAD
    g = torch.Generator().manual_seed(seed)
    lim = (6.0 / W["d"]) ** 0.5
    ent = (torch.rand((W["n_ent"], W["d"]), generator=g) * 2 - 1) * lim
    rel = (torch.rand((W["n_rel"], W["d"]), generator=g) * 2 - 1) * lim
    return {"ent_emb": ent, "rel_inter": rel}


def make_engine(device):
    from KGE._b200 import engine as E, spec
    W = WORKLOAD

    class _BCE:  # descriptor only: spec maps it by class name
        pass
    _BCE.__name__ = "BinaryCrossEntropyLoss"
    cfg = spec.make_cfg("DistMult", W["n_ent"], W["n_rel"], W["d"], loss_fn=_BCE(), negative_ratio=W["k"],
                        corrupt_side=W["side"], constraint=True, constraint_weight=1.0)
    T = init_tables()
    return E.Engine(cfg, {spec.SLOTS["DistMult"][k]: v for k, v in T.items()}, device=device)


def algorithmic_bytes_step():
    """forward gather 4*((2+k')*d + d) B per positive + 3 ids (k_step's share);
    row update 6*4*U*row_floats with U ~ all rows touched at this k (k_update's share)."""
    W = WORKLOAD
    k_eff = 2 * (W["k"] // 2)
    fwd = W["B"] * (4 * ((2 + k_eff) * W["d"] + W["d"]) + 3 * 8)
    upd = 6 * 4 * (W["n_ent"] + W["n_rel"]) * W["d"]
    return fwd, upd


def run_ours(args, rank, world, device):
    W = WORKLOAD
    B, k_eff = W["B"], 2 * (W["k"] // 2)
    eng = make_engine(device)
    eng.dp_w_private = world > 1 and os.environ.get("KGE_DP_W_PRIVATE", "1") != "0"   # only wn travels per step (kge_grads.w_private)
    train = zipf_triples(W["n_train"], W["n_ent"], W["n_rel"], 12345)
    # data-parallel (world > 1): one GLOBAL batch of world*B positives per step, identical on every
    # rank; rank r owns positives [r*B, (r+1)*B) (weak scaling: per-GPU batch fixed).  Gradients are
    # exchanged inside Engine.train_step_dp (peer staging over NVLink, or one NCCL all-reduce).
    rng = np.random.default_rng(100)
    n_batches = 64
    host_batches = [torch.from_numpy(train[rng.integers(0, len(train), B * world)][rank * B:(rank + 1) * B].copy())
                    .pin_memory() for _ in range(n_batches)]
    dev_batches = [b.to(device) for b in host_batches]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)      # > 126 MB L2
    loss_dev = torch.zeros(1, dtype=torch.float32, device=device)
    seed = 12345
    step_no = [0]

    def one_step(batch, nxt=None):
        """One step through the plain C-ABI call; nxt = the following batch (look-ahead grouping, single GPU)."""
        if world > 1:
            eng.train_step_dp(batch, seed, step_no[0], B * world, rank * B, loss_out=loss_dev, next_triples=nxt)
        else:
            eng.train_step(batch, seed, step_no[0], is_train=True, loss_out=loss_dev, next_triples=nxt)
        step_no[0] += 1

    graph = None
    if os.environ.get("KGE_BENCH_GRAPH", "1") != "0":
        graph = eng.step_graph(B, seed) if world == 1 else eng.step_graph_dp(B, seed, B * world, rank * B)

    def graph_step(batch, nxt, host_loss=None):
        """The same step through the captured graph (kge_step_graph): batch / nxt from device or pinned host memory."""
        graph.launch(batch, step_no[0], next_triples=nxt, host_loss=host_loss)
        step_no[0] += 1

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    clk = Clocks(torch.cuda.current_device()).start()
    for i in range(args.warmup):
        one_step(dev_batches[i % n_batches], dev_batches[(i + 1) % n_batches])
    if graph is not None:
        for i in range(args.warmup):
            graph_step(dev_batches[i % n_batches], dev_batches[(i + 1) % n_batches])
    barrier()

    # ---- value: inputs resident in HBM, L2 flushed between steps, per-step CUDA events.  Single GPU: the step is
    # the captured graph with the next batch's grouping under it (the product's training loop); the plain-call
    # figure (same kernels, ~20 launches per step issued by the host) is kept beside it. ----
    def timed_cold(step_fn):
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        barrier()
        for i in range(args.steps):
            flush.zero_()
            ev[i][0].record()
            step_fn(dev_batches[i % n_batches], dev_batches[(i + 1) % n_batches])
            ev[i][1].record()
        barrier()
        return sum(a.elapsed_time(b) for a, b in ev)
    ms_cold_plain = timed_cold(one_step)
    ms_cold = timed_cold(graph_step) if graph is not None else ms_cold_plain

    # ---- the same loop again with the library's per-phase events on (they sit between the
    # kernels, so this pass is for the phase breakdown / roofline only, not for `value`) ----
    n_prof = min(args.steps, 100)
    eng.profile_begin(n_prof)
    barrier()
    for i in range(n_prof):
        flush.zero_()
        one_step(dev_batches[i % n_batches], dev_batches[(i + 1) % n_batches])
    barrier()
    phases, n_calls = eng.profile_end()

    # ---- the graph's own timeline: a second graph created with phase events inside (kge_step_graph_phases) ----
    graph_phases = None
    if graph is not None:
        from KGE._b200 import lib as _L
        eng.set_option(_L.OPT_GRAPH_PROFILE, 1)
        gprof = eng.step_graph(B, seed) if world == 1 else eng.step_graph_dp(B, seed, B * world, rank * B)
        eng.set_option(_L.OPT_GRAPH_PROFILE, 0)
        acc, tot, n_gp = {}, 0.0, min(args.steps, 40)
        for i in range(n_gp + 4):
            flush.zero_()
            gprof.launch(dev_batches[i % n_batches], step_no[0], next_triples=dev_batches[(i + 1) % n_batches])
            step_no[0] += 1
            ph, t_all = gprof.phases()
            if i >= 4:
                tot += t_all
                for kk, v in ph.items():
                    acc[kk] = acc.get(kk, 0.0) + v
        graph_phases = {kk: round(v / n_gp, 5) for kk, v in acc.items() if v > 0}
        graph_phases["first_to_last_node"] = round(tot / n_gp, 5)
        del gprof
        barrier()

    # ---- steady state (no flush): what a training loop sees, state is L2 resident ----
    barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(args.steps):
        (graph_step if graph is not None else one_step)(dev_batches[i % n_batches], dev_batches[(i + 1) % n_batches])
    b.record()
    barrier()
    ms_warm = a.elapsed_time(b)

    # ---- e2e: pinned host batch -> H2D -> C-ABI step -> D2H of the loss, every step ----
    loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()
    stage = torch.empty((B, 3), dtype=torch.int64, device=device)
    barrier()
    seen_e2e = 0.0
    t0 = time.perf_counter()
    for i in range(args.steps):
        if graph is None and world > 1:
            stage.copy_(host_batches[i % n_batches], non_blocking=True)
            one_step(stage)
            loss_host.copy_(loss_dev, non_blocking=True)
        elif graph is not None:   # host buffers through the captured graph: H2D of this (first step) / the next batch,
            graph_step(host_batches[i % n_batches], host_batches[(i + 1) % n_batches], loss_host)   # step, D2H of the loss
            graph.wait_loss()     # the host has THIS step's loss before it issues the next one (the reference reads
            seen_e2e += float(loss_host[0])                                    # loss.numpy() per batch); one GPU: the
            continue              # graph stores it into the pinned word as soon as it is final (kge_step_graph_wait_loss)
        else:           # host buffers straight through the C-ABI: H2D + step + D2H enqueued by kge_train_step_host
            eng.train_step_host(host_batches[i % n_batches], seed, step_no[0], loss_host)
            step_no[0] += 1
        torch.cuda.current_stream().synchronize()      # the reference reads loss.numpy() per batch
        seen_e2e += float(loss_host[0])
    barrier()                     # every step's kernels have finished inside the timed region
    ms_e2e = (time.perf_counter() - t0) * 1e3
    assert math.isfinite(seen_e2e), "a loss read in the end-to-end loop was not finite"
    if graph is not None:         # what the host read last is the last step's loss, bit for bit
        assert float(loss_host[0]) == float(graph.loss.item()), "host loss differs from the device loss of the last step"
    ms_e2e_plain = None
    if graph is not None and world == 1:       # the same through kge_train_step_host (no graph, no look-ahead), for comparison
        barrier()
        t0 = time.perf_counter()
        for i in range(args.steps):
            eng.train_step_host(host_batches[i % n_batches], seed, step_no[0], loss_host)
            step_no[0] += 1
            torch.cuda.current_stream().synchronize()
        barrier()
        ms_e2e_plain = (time.perf_counter() - t0) * 1e3

    # ---- the same with the loss of step i read one step late (what KGEModel.train does even more
    # lazily: it keeps the per-batch losses on the device and syncs once per epoch); extra information,
    # `e2e` stays the per-step-synchronised figure ----
    loss2 = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(2)]
    done = [torch.cuda.Event(), torch.cuda.Event()]
    seen = 0.0
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        if graph is None and world > 1:
            stage.copy_(host_batches[i % n_batches], non_blocking=True)
            one_step(stage)
            loss2[i & 1].copy_(loss_dev, non_blocking=True)
        elif graph is not None:
            graph_step(host_batches[i % n_batches], host_batches[(i + 1) % n_batches], loss2[i & 1])
        else:
            eng.train_step_host(host_batches[i % n_batches], seed, step_no[0], loss2[i & 1])
            step_no[0] += 1
        done[i & 1].record()
        if i > 0:
            done[(i - 1) & 1].synchronize()
            seen += float(loss2[(i - 1) & 1][0])
    torch.cuda.current_stream().synchronize()
    barrier()
    ms_e2e_lag = (time.perf_counter() - t0) * 1e3

    # clocks: at least two samples under the bench's own load (same step, untimed) — every rank keeps
    # stepping for the same number of rounds so data-parallel collectives stay matched
    need = torch.tensor([1 if (clk.proc is not None and len(clk.rows) < 2) else 0], device=device)
    for _ in range(60):
        if world > 1:
            torch.distributed.all_reduce(need, op=torch.distributed.ReduceOp.MAX)
        if int(need.item()) == 0:
            break
        for i in range(200):
            one_step(dev_batches[i % n_batches], dev_batches[(i + 1) % n_batches])
        torch.cuda.synchronize()
        need[0] = 1 if (clk.proc is not None and len(clk.rows) < 2) else 0
    clk.stop()

    eng.dp_sync_weights()                 # w of the rows other ranks own (no-op unless dp_w_private)
    t = torch.tensor([ms_cold, ms_warm, ms_e2e, ms_e2e_lag], dtype=torch.float64, device=device)
    if world > 1:
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    ms_cold, ms_warm, ms_e2e, ms_e2e_lag = t.tolist()
    triples = world * B * (1 + k_eff) * args.steps
    return {"eng": eng, "graph_phases": graph_phases, "ms_cold_plain": ms_cold_plain, "ms_e2e_plain": ms_e2e_plain, "graph": graph is not None, "ms_cold": ms_cold, "ms_warm": ms_warm, "ms_e2e": ms_e2e, "ms_e2e_lag": ms_e2e_lag, "triples": triples,
            "phases": phases, "n_calls": n_calls, "n_prof": n_prof, "clocks": clk.summary(), "train": train,
            "loss": float(loss_host.item())}


def run_eval(eng, train, rank, world, device, reps=3):
    """Filtered ranking, both sides, Q queries per side.  DistMult scores are dot products: kge_rank runs the
    tcgen05 GEMM path (csrc/rank_dot_tc.cuh).  One GPU: the whole entity table.  N > 1: every rank scores ALL
    queries against its slice of the entity table (kge_rank ent_lo / ent_hi) and the int32 counts are summed by one
    NCCL all-reduce — the collective is inside both timed regions.

    kernel-only (`ms`): queries, filter index and ranges resident; kge_rank (+ all-reduce) only.
    e2e (`ms_e2e`): what evaluate() does per call, from HOST buffers: H2D of positive_X and of the queries,
    kge_filter_index_build (one sort per side), kge_filter_lookup, kge_rank, all-reduce, D2H of the counts."""
    from KGE._b200 import dp as _dp
    W = WORKLOAD
    Q = W["eval_Q"]
    rng = np.random.default_rng(777)
    q_all = train[rng.integers(0, len(train), Q)]
    e_lo, e_hi = _dp.split_even(W["n_ent"], world, rank)
    qd = torch.from_numpy(q_all).to(device)
    index = {s: eng.filter_index(train, s) for s in ("h", "t")}
    filt = {s: index[s].lookup(qd) for s in ("h", "t")}
    side_id = {"h": 0, "t": 1}

    def counts(s, q, f):
        c = eng.rank_counts(q, side_id[s], f[0], f[1], ent_lo=e_lo, ent_hi=e_hi)
        if world > 1:
            torch.distributed.all_reduce(c)
        return c
    cnts = {s: counts(s, qd, filt[s]) for s in ("h", "t")}
    torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    eng.profile_begin(2 * reps)
    a.record()
    for _ in range(reps):
        for s in ("h", "t"):
            counts(s, qd, filt[s])
    b.record()
    torch.cuda.synchronize()
    phases, n_calls = eng.profile_end()

    # ---- e2e: host positive_X + host queries -> ranks on the host, one evaluate() call per side ----
    # KGEModel.evaluate (BaseModel._rank_block) uploads positive_X and the queries, builds the filter index ONCE and
    # ranks the queries in blocks of Q; the call timed here has the FB15k test split's size (59 071 queries = 4 blocks)
    Qe, blk = W["eval_Q_e2e"], Q
    q_e2e = np.concatenate([q_all, train[rng.integers(0, len(train), Qe - Q)]])
    pos_host = torch.from_numpy(train).pin_memory()
    q_host = torch.from_numpy(q_e2e).pin_memory()
    out_host = {s: torch.zeros(Qe, dtype=torch.int32).pin_memory() for s in ("h", "t")}

    def evaluate_call(s, marks=None):
        def mark(name):
            if marks is not None:
                e = torch.cuda.Event(enable_timing=True)
                e.record()
                marks.append((name, e))
        mark("start")
        pos_dev = pos_host.to(device, non_blocking=True)
        q_dev = q_host.to(device, non_blocking=True)
        mark("h2d")
        idx = eng.filter_index(pos_dev, s)
        mark("index_build")
        for lo in range(0, Qe, blk):
            qb = q_dev[lo:lo + blk]
            f = idx.lookup(qb)
            mark("lookup")
            c = counts(s, qb, f)
            mark("rank")
            out_host[s][lo:lo + blk].copy_(c, non_blocking=True)
            mark("d2h")
    for s in ("h", "t"):          # one untimed call per side (first cudaMalloc of the staging / index / ragged-block buffers)
        evaluate_call(s)
    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        for s in ("h", "t"):
            evaluate_call(s)
        torch.cuda.synchronize()
    ms_e2e = (time.perf_counter() - t0) * 1e3
    for s in ("h", "t"):
        assert torch.equal(out_host[s][:Q], cnts[s].cpu()), "e2e ranks differ from the resident-input ranks"
    seg = {}                          # where one call's device time goes (an extra, untimed call with events between the stages)
    marks = []
    evaluate_call("t", marks)
    torch.cuda.synchronize()
    for (_, e0), (name, e1) in zip(marks[:-1], marks[1:]):
        seg[name] = round(seg.get(name, 0.0) + e0.elapsed_time(e1), 4)

    t = torch.tensor([a.elapsed_time(b), ms_e2e], dtype=torch.float64, device=device)
    if world > 1:
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    ms, ms_e2e = t.tolist()
    ranks = {s: (c.cpu().numpy() + 1).astype(np.float64) for s, c in cnts.items()}
    both = np.concatenate([ranks["h"], ranks["t"]])
    gemm_ms = phases["rank_gemm"] / max(n_calls, 1)
    flops = 2.0 * (e_hi - e_lo) * W["d"] * Q                  # algorithmic, this rank's slice: 2*N*d per query
    return {"queries": 2 * Q * reps, "queries_e2e": 2 * Qe * reps, "e2e_segments_ms": seg, "ms": ms, "ms_e2e": ms_e2e, "kernel_ms": phases["rank"], "gemm_ms": gemm_ms,
            "h2d_bytes_per_call": int(pos_host.numel() * 8 + q_host.numel() * 8), "d2h_bytes_per_call": 4 * Qe,
            "fix_ms": phases["rank_fix"] / max(n_calls, 1), "gemm_tflops": flops / (gemm_ms * 1e-3) / 1e12 if gemm_ms else None,
            "gemm_flops": flops, "sharding": "entity shards + NCCL all-reduce of int32 counts" if world > 1 else "none",
            "mrr": float(np.mean(1 / both)), "hits10": float(np.mean(both <= 10)), "ranks": ranks, "queries_used": q_all}


def cpu_oracle_eval(eng, train, ev, n_queries, threads):
    """The reference's get_rank loop (BaseModel.py:602-654: per query, two scans of positive_X, score all
    entities) on host cores, on the first n_queries of the GPU's query set.  GPU ranks must equal the oracle's
    except where an ELIGIBLE candidate (not the true entity, not filtered) scores within 1e-5 relative of the
    true triple (north_star's documented tie band): `ranks_outside_tie_band` must be 0."""
    from oracle import ref_ops as O, ref_models as RM, ref_step as RS
    from KGE._b200 import spec
    W = WORKLOAD
    torch.set_num_threads(threads)
    ref = RM.DistMult(W["n_ent"], W["n_rel"], W["d"], loss_fn=O.BinaryCrossEntropyLoss())
    Wt = RM.Weights({k: eng.tab[s].view.detach().cpu() for k, s in spec.SLOTS["DistMult"].items()})
    qs = ev["queries_used"][:n_queries]
    t0 = time.perf_counter()
    r_cpu = {s: np.array([RS.get_rank(ref, Wt, x, train, s) for x in qs], dtype=np.float64) for s in ("h", "t")}
    dt = time.perf_counter() - t0
    cpu_all = np.concatenate([r_cpu["h"], r_cpu["t"]])
    gpu_all = np.concatenate([ev["ranks"]["h"][:n_queries], ev["ranks"]["t"][:n_queries]])
    in_band = outside = 0
    for s, col, fs in (("h", 0, 2), ("t", 2, 0)):
        for q, x in enumerate(qs):
            got, want = int(ev["ranks"][s][q]), int(r_cpu[s][q])
            sc = RS.get_scores_all(ref, Wt, x, s).double().numpy()
            with torch.no_grad():
                one = lambda v: torch.tensor([int(v)])                    # noqa: E731
                p = float(ref.score_hrt(Wt, one(x[0]), one(x[1]), one(x[2]))[0])
            ex = train[(train[:, 1] == x[1]) & (train[:, fs] == x[fs]), col]
            elig = np.ones(len(sc), bool)
            elig[ex] = False
            true_filtered = bool((ex == x[col]).any())
            elig[x[col]] = False
            want -= int((not true_filtered) and sc[x[col]] > p)           # the reference's self-count
            if got == want:
                continue
            near = int((np.abs(sc - p) <= 1e-5 * max(1.0, abs(p)))[elig].sum())
            if abs(got - want) <= near:
                in_band += 1
            else:
                outside += 1
    return {"value": 2 * len(qs) / dt, "unit": "queries/s", "cores": threads, "kind": "port",
            "sample": f"{2 * len(qs)} queries (both sides of the first {len(qs)} eval triples), filter = train set",
            "mrr_cpu": float(np.mean(1 / cpu_all)), "mrr_gpu_same_queries": float(np.mean(1 / gpu_all)),
            "hits10_cpu": float(np.mean(cpu_all <= 10)), "hits10_gpu_same_queries": float(np.mean(gpu_all <= 10)),
            "ranks_identical": int((cpu_all == gpu_all).sum()), "ranks_compared": int(len(cpu_all)),
            "ranks_inside_tie_band": in_band, "ranks_outside_tie_band": outside}


def cpu_oracle_steps(n_steps, threads, warmup=1):
    """The reference math on host cores: oracle/ref_step.run_single_batch on the bench workload
    (`warmup` untimed steps first)."""
    from oracle import ref_ops as O, ref_models as RM, ref_step as RS
    W = WORKLOAD
    torch.set_num_threads(threads)
    ref = RM.DistMult(W["n_ent"], W["n_rel"], W["d"], loss_fn=O.BinaryCrossEntropyLoss())
    Wt = RM.Weights(init_tables())
    opt = RS.TFAdam()
    train = zipf_triples(20000, W["n_ent"], W["n_rel"], 12345)
    rng = np.random.default_rng(5)
    times = []
    for step in range(n_steps + warmup):
        X = train[rng.integers(0, len(train), W["B"])]
        t0 = time.perf_counter()
        ids, ke = RS.sample_negatives(12345, step, X, W["k"], W["side"], W["n_ent"])
        RS.run_single_batch(ref, Wt, opt, X, ids, ke, W["side"], True)
        if step >= warmup:
            times.append(time.perf_counter() - t0)
    k_eff = 2 * (W["k"] // 2)
    return W["B"] * (1 + k_eff) * len(times) / sum(times), sum(times) / len(times)


# ---------------------------------------------------------------------------------------------------
# C5 (BASELINE.json configs[4]): synthetic 5 M entities x 1 k relations, TransE d = 128, L1, hinge, uniform
# 'h+t' k = 32 — the one configuration whose tables (2.56 GB + Adam slots) do not sit in L2.
#   N = 1 : row-lazy step on one GPU (kge_tables.lazy_t);   N > 1 : entity tables SHARDED over the ranks
#   (kge_shard_*, the class KGEModel.train uses: _b200.shard.ShardedTrainer), weak scaling: 8192 positives per rank.
# ---------------------------------------------------------------------------------------------------
C5 = {"workload": "C5 TransE d=128 L1 hinge uniform h+t k=32, N=5,000,000 R=1,000 synthetic (h,t uniform, r Zipf)",
      "n_ent": 5_000_000, "n_rel": 1000, "d": 128, "k": 32, "B_local": 8192, "eval_Q": 10000}


def c5_batches(n, B, seed):
    W = C5
    rng = np.random.default_rng(seed)
    wrel = 1.0 / np.arange(1, W["n_rel"] + 1)
    wrel /= wrel.sum()
    return [np.stack([rng.integers(0, W["n_ent"], B), rng.choice(W["n_rel"], size=B, p=wrel),
                      rng.integers(0, W["n_ent"], B)], 1).astype(np.int64) for _ in range(n)]


def run_c5(args, rank, world, device, pk):
    from KGE._b200 import engine as E, spec, shard as SH, dp as _dp
    W = C5
    N, R, d, k = W["n_ent"], W["n_rel"], W["d"], W["k"]
    B_local, B_global = W["B_local"], W["B_local"] * world
    steps, warm = max(10, min(args.steps, 60)), 5

    class _D:
        pass

    def desc(name, **kw):
        o = type(name, (), {})()
        for kk, v in kw.items():
            setattr(o, kk, v)
        return o
    cfg = spec.make_cfg("TransE", N, R, d, score_fn=desc("LpDistance", p=1), loss_fn=desc("PairwiseHingeLoss", margin=1.0),
                        negative_ratio=k, corrupt_side="h+t", constraint=True)
    g = torch.Generator(device=device).manual_seed(12345)          # same stream on every rank (same GPU model)
    lim = 6.0 / d ** 0.5
    tabs = {0: (torch.rand((N, d), generator=g, device=device) * 2 - 1) * lim,
            2: (torch.rand((R, d), generator=g, device=device) * 2 - 1) * lim}
    batches = c5_batches(8, B_global, 4242)                       # GLOBAL batches, identical on every rank
    dev_batches = [torch.from_numpy(b).to(device) for b in batches]
    loss = torch.zeros(1, dtype=torch.float32, device=device)
    parity = None
    if world == 1:
        eng = E.Engine(cfg, tabs, device=device, lazy=True)
        step_fn = lambda X, i: eng.train_step(X, 3, i, loss_out=loss)                  # noqa: E731
        prof_eng, slots_per_step = eng, 1
    else:
        tr = SH.ShardedTrainer(cfg, tabs, B_global, device)
        step_fn = lambda X, i: tr.train_step(X, 3, i, loss_out=loss)                   # noqa: E731
        prof_eng, slots_per_step = tr.ranks[0].eng, 3
        # parity: the sharded step equals a one-GPU row-lazy step on the whole global batch (rank 0 runs it)
        single = E.Engine(cfg, tabs, device=device, lazy=True) if rank == 0 else None
        ls, lm = [], []
        for i in range(2):
            lm.append(float(step_fn(dev_batches[i], i).item()))
            if single is not None:
                ls.append(float(single.train_step(dev_batches[i], 3, i).item()))
        if rank == 0:
            parity = {"loss_sharded": lm, "loss_single_gpu_row_lazy": ls,
                      "max_rel_diff": max(abs(a - b) / max(1.0, abs(b)) for a, b in zip(lm, ls)), "tolerance": 1e-5}
        del single
        torch.cuda.empty_cache()
    del tabs

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()
    for i in range(warm):
        step_fn(dev_batches[i % 8], 10 + i)
    barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(steps):
        step_fn(dev_batches[i % 8], 100 + i)
    b.record()
    barrier()
    ms = a.elapsed_time(b)
    n_prof = min(steps, 20)
    prof_eng.profile_begin(n_prof * slots_per_step)
    for i in range(n_prof):
        step_fn(dev_batches[i % 8], 1000 + i)
    barrier()
    phases, n_calls = prof_eng.profile_end()
    # e2e: pinned host GLOBAL batch -> H2D -> step -> D2H of the loss, stream synchronised every step
    host_batches = [torch.from_numpy(x).pin_memory() for x in batches]
    loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()
    stage = torch.empty((B_global, 3), dtype=torch.int64, device=device)
    barrier()
    t0 = time.perf_counter()
    for i in range(steps):
        stage.copy_(host_batches[i % 8], non_blocking=True)
        step_fn(stage, 2000 + i)
        loss_host.copy_(loss, non_blocking=True)
        torch.cuda.current_stream().synchronize()
    barrier()
    ms_e2e = (time.perf_counter() - t0) * 1e3

    # filtered eval: Q queries x 2 sides against all 5 M entities (CUDA-core Lp tile kernel), entity shards +
    # all-reduce of the counts at N > 1 (the gathered table is replicated for evaluation only)
    ev_eng = eng if world == 1 else tr.eval_engine()
    Qn = W["eval_Q"]
    known = np.concatenate(c5_batches(4, 50_000, 99))
    rngq = np.random.default_rng(5)
    Xq = torch.from_numpy(known[rngq.integers(0, len(known), Qn)]).to(device)
    e_lo, e_hi = _dp.split_even(N, world, rank)
    idx = {s: ev_eng.filter_index(known, s) for s in ("h", "t")}

    def rank_all():
        out = []
        for s, sid in (("h", 0), ("t", 1)):
            c = ev_eng.rank_counts(Xq, sid, *idx[s].lookup(Xq), ent_lo=e_lo, ent_hi=e_hi)
            if world > 1:
                torch.distributed.all_reduce(c)
            out.append(c)
        return out
    ev_eng.rank_counts(Xq[:256], 1)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    cnt = rank_all()
    e1.record()
    barrier()
    ms_eval = e0.elapsed_time(e1)
    t = torch.tensor([ms, ms_e2e, ms_eval], dtype=torch.float64, device=device)
    if world > 1:
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    ms, ms_e2e, ms_eval = t.tolist()
    if rank != 0:
        return None
    k_eff = 2 * (k // 2)
    triples = B_global * (1 + k_eff) * steps
    alg = B_global * (7 * 4 * ((2 + k_eff) * d + d) + 24)           # 125 452 B per positive
    gbs_per_gpu = alg / world / (ms / steps * 1e-3) / 1e9
    rk = np.concatenate([c.cpu().numpy() for c in cnt]).astype(np.float64) + 1
    out = {"workload": W["workload"], "metric": METRIC, "value": triples / (ms * 1e-3), "unit": "triples/s",
           "n_gpus": world, "global_batch": B_global, "per_gpu_batch": B_local, "steps": steps, "ms_per_step": ms / steps,
           "tables": "row-lazy, one GPU (kge_tables.lazy_t)" if world == 1 else
                     "entity tables sharded over the ranks (kge_shard_*), relation table replicated",
           "e2e": {"value": triples / (ms_e2e * 1e-3), "unit": "triples/s", "h2d_bytes_per_step": B_global * 24,
                   "d2h_bytes_per_step": 4},
           "roofline": {"bound": "hbm", "achieved": gbs_per_gpu, "peak": pk["hbm_gbs"], "unit": "GB/s per GPU",
                        "frac": gbs_per_gpu / pk["hbm_gbs"], "algorithmic_bytes_per_step": alg,
                        "note": "whole step (lookup, replay, gather, scores, backward, row update), 3 802 B per triple"},
           "phase_ms_per_step": {kk: round(v / max(n_prof, 1), 5) for kk, v in phases.items() if v > 0},
           "profile_slots": n_calls,
           "nvlink_bytes_per_rank_per_step": 0 if world == 1 else int(
               (world - 1) / world * B_local * (2 + k_eff) * (2 * d * 4 + 4)),
           "nvlink_note": "owner-applies: one row read + one gradient row written per looked-up remote row, + its id",
           "parity": parity,
           "eval": {"metric": "filtered-eval queries/sec", "value": 2 * Qn / (ms_eval * 1e-3), "unit": "queries/s",
                    "queries": 2 * Qn, "ms": ms_eval, "kernel": "k_rank_lp_tile (CUDA-core L1 tile)",
                    "sharding": "entity shards + NCCL all-reduce of int32 counts" if world > 1 else "none",
                    "lp_ops_per_s_T": 2 * Qn * N * d * 3 / (ms_eval * 1e-3) / 1e12,
                    "mrr": float(np.mean(1 / rk)), "hits10": float(np.mean(rk <= 10))},
           "final_loss": float(loss_host.item())}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-steps", type=int, default=30)
    ap.add_argument("--cpu-queries", type=int, default=64)
    ap.add_argument("--no-c5", action="store_true", help="skip the C5 (5 M entities) leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    W = WORKLOAD
    k_eff = 2 * (W["k"] // 2)
    cores = os.cpu_count() or 1
    config = {"workload": W["workload"], "global_batch": W["B"] * world, "negatives_per_positive": k_eff,
              "optimizer": "Adam(tf defaults), clip_by_norm 5", "l2": "flushed between timed steps (256 MiB memset)",
              "parallelism": f"dp{world}"}

    if args.impl == "reference":
        # the reference's own CPU path for this step (TensorFlow 2.5 is not installable here: the oracle port),
        # all host threads, rank 0 only; each timed step is one full step of the workload (B = 1024, 65 triples
        # per positive), at most --cpu-steps of them so the run ends within minutes
        if rank != 0:
            return
        steps = min(args.steps, args.cpu_steps)
        tps, spstep = cpu_oracle_steps(steps, cores, warmup=args.warmup)
        sample = (f"{steps} full steps of the bench workload (B={W['B']}, k'={k_eff}) after {args.warmup} warm-up steps"
                  + (f" (--steps {args.steps} capped at --cpu-steps {args.cpu_steps})" if steps < args.steps else ""))
        print(json.dumps({"impl": "reference", "metric": METRIC, "value": tps, "unit": "triples/s", "n_gpus": args.gpus,
                          "steps": steps, "warmup": args.warmup, "ms_per_step": spstep * 1e3, "higher_is_better": True,
                          "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
                          "cpu_baseline": {"value": tps, "unit": "triples/s", "cores": cores, "kind": "port",
                                           "sample": sample},
                          "e2e": {"value": tps, "unit": "triples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                          "note": "reference TF path unavailable on this box (TensorFlow 2.5 not installable); "
                                  "this is the CPU oracle port of the same step"}))
        return

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=device)
    r = run_ours(args, rank, world, device)
    ev = run_eval(r["eng"], r["train"], rank, world, device)
    c5 = None
    if not args.no_c5:
        eng_keep = r.pop("eng")
        c5 = run_c5(args, rank, world, device, peaks())
        r["eng"] = eng_keep
    if rank != 0:
        if world > 1:
            torch.distributed.destroy_process_group()
        return
    pk = peaks()
    fwd_bytes, upd_bytes = algorithmic_bytes_step()
    step_ms = r["phases"]["step"] / max(r["n_calls"], 1)
    achieved = fwd_bytes / (step_ms * 1e-3) / 1e9
    # per-step milliseconds of each phase (library-side CUDA events); "sort" (keys + radix sort +
    # segment table) runs on the side stream concurrently with constraint + step
    share = {k: round(v / max(r["n_calls"], 1), 5) for k, v in r["phases"].items() if v > 0}
    cpu_tps, cpu_s = cpu_oracle_steps(args.cpu_steps, cores)
    assert r["n_calls"] == r["n_prof"], ("profile slots", r["n_calls"], r["n_prof"])     # one slot per step, also data-parallel
    cpu_ev = cpu_oracle_eval(r["eng"], r["train"], ev, args.cpu_queries, cores)
    out = {
        "metric": METRIC, "value": r["triples"] / (r["ms_cold"] * 1e-3), "unit": "triples/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_cold"] / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config,
        "exchange": (getattr(r["eng"], "_dp_kind", None) or "none") if world > 1 else "none",
        "exchange_w_private": bool(world > 1 and getattr(getattr(r["eng"], "_peer", None), "w_private", False)),
        "exchange_multicast": bool(world > 1 and getattr(getattr(getattr(r["eng"], "_peer", None), "ptabs", None), "mc_ptr", 0)),
        "clocks": r["clocks"],
        "value_steady_state_l2_warm": r["triples"] / (r["ms_warm"] * 1e-3),
        "step_api": (("kge_step_graph_launch_dp" if world > 1 else "kge_step_graph_launch") +
                     " (captured CUDA graph, next batch grouped under the step)" if r["graph"] else
                     "kge_train_step_dp_next" if world > 1 else "kge_train_step_next"),
        "graph_phase_ms_per_step": r["graph_phases"],
        "graph_phase_note": "events recorded BY a second, instrumented graph (kge_step_graph_phases); 'sort' is the NEXT batch's grouping chain, concurrent with step + update",
        "value_plain_calls": r["triples"] / (r["ms_cold_plain"] * 1e-3),
        "value_plain_calls_note": "same kernels through kge_train_step_next: ~20 launches / event calls per step issued by the host",
        "e2e_plain_calls": (r["triples"] / (r["ms_e2e_plain"] * 1e-3)) if r["ms_e2e_plain"] else None,
        "e2e": {"value": r["triples"] / (r["ms_e2e"] * 1e-3), "unit": "triples/s",
                "h2d_bytes_per_step": W["B"] * 3 * 8, "d2h_bytes_per_step": 4,
                "sync": ("the host waits for and reads every step's loss before it issues the next step "
                         "(kge_step_graph_wait_loss: one GPU — the graph reads the pinned batch and stores the loss "
                         "into the pinned word itself, the wait ends when the loss is final, the update kernels of "
                         "step i overlap the launch of step i+1; N > 1 — stream synchronised); all kernels of all "
                         "steps complete inside the timed region; L2 is NOT flushed between these steps (a flush kernel would sit "
                         "inside the wall-clock region): compare with value_steady_state_l2_warm, not with the cold-L2 value") if r["graph"] else
                        "stream synchronised and the loss read on the host after every step"},
        "e2e_loss_one_step_late": {"value": r["triples"] / (r["ms_e2e_lag"] * 1e-3), "unit": "triples/s",
                                   "note": "same H2D / step / D2H per step, loss of step i read while step i+1 runs "
                                           "(KGEModel.train reads losses once per epoch)"},
        "gpu_launches": args.steps * (9 if world == 1 else 13),      # single GPU: k_set_dyn, k_stage_next, k_step, k_keys, k_scan, k_scatter, k_canon, k_update_heavy, k_update_fast
        "gpu_launches_note": "own kernels per step: k_constrain, k_step (loss/clip reduction folded into its last warp), "
        "k_update_fast on the caller's stream; k_keys, k_scan, k_scatter, k_canon, k_update_heavy on the side stream "
        "(+ one memset there); data-parallel runs split the update pair into k_update_fast/heavy<3> (row sums stored "
        "into every rank's staging slot over NVLink) + k_push_scal, one cross-GPU barrier, then k_dp_scales + "
        "k_update_fast<4> (config.exchange = peer), or the same around one NCCL all-reduce (exchange = nccl); exchange = owner (12 kernels: k_step, k_keys, k_scan, k_scatter, k_canon, k_update_heavy<3>, k_update_fast<3>, k_push_scal, k_peer_barrier, k_dp_scales, k_update_fast<4>, k_peer_barrier): row sums go to the row owner only, the owner runs k_update_fast<4> on its rows and stores the new rows into every replica over NVLink, all enqueued by one kge_train_step_dp call",
        "roofline": {"kernel": "k_step<MDistMult,2,true>", "bound": "hbm", "achieved": achieved, "peak": pk["hbm_gbs"],
                     "unit": "GB/s", "frac": achieved / pk["hbm_gbs"], "traffic": ncu_traffic("k_step<MDistMult"),
                     "traffic_note": "ncu dram read+write per launch (cold caches); below the algorithmic bytes because "
                                     "the 12 MB entity table is read from HBM once and re-gathered from L2",
                     "peak_src": pk["src"],
                     "algorithmic_bytes_per_launch": fwd_bytes, "avg_launch_ms": step_ms, "phase_ms_per_step": share},
        "roofline_all": roofline_all(share, fwd_bytes, upd_bytes, pk, world),
        "cpu_baseline": {"value": cpu_tps, "unit": "triples/s", "cores": cores, "kind": "port",
                         "sample": f"{args.cpu_steps} full steps of the same workload on the host (oracle port; "
                                   "TensorFlow reference not installable)"},
        "eval": {"metric": "filtered-eval queries/sec", "value": ev["queries"] / (ev["ms"] * 1e-3), "unit": "queries/s",
                 "value_note": "kernel-only: queries, filter index and ranges resident; kge_rank (+ all-reduce of counts at N > 1)",
                 "e2e": {"value": ev["queries_e2e"] / (ev["ms_e2e"] * 1e-3), "unit": "queries/s", "queries": ev["queries_e2e"], "device_ms_per_call": ev["e2e_segments_ms"],
                         "h2d_bytes_per_call": ev["h2d_bytes_per_call"], "d2h_bytes_per_call": ev["d2h_bytes_per_call"],
                         "what": "one evaluate() call per side over 59 071 queries (the FB15k test split): H2D of positive_X "
                                 "(483 142 triples) and of the queries, kge_filter_index_build once (one 64-bit radix sort), then "
                                 "per block of 16 384 queries kge_filter_lookup, kge_rank, all-reduce, D2H of the counts; host "
                                 "synchronised after both sides; one untimed warm-up call per side, then 3 timed"},
                 "queries": ev["queries"], "sharding": ev["sharding"],
                 "kernel": "k_rank_dot_tc (tcgen05 kind::tf32) + k_rank_dot_fix/filter",
                 "mrr": ev["mrr"], "hits10": ev["hits10"],
                 "roofline": {"bound": "tensor", "achieved": ev["gemm_tflops"], "peak": pk["tf32_tflops"],
                              "unit": "TFLOP/s", "frac": (ev["gemm_tflops"] or 0) / pk["tf32_tflops"],
                              "peak_note": pk["tf32_src"],
                              "algorithmic_flops_per_launch": ev["gemm_flops"],
                              "avg_launch_ms": ev["gemm_ms"], "fixup_filter_ms": ev["fix_ms"], "traffic": ncu_traffic("k_rank_dot_tc")},
                 "cpu_baseline": cpu_ev},
        "c5": c5,
        "final_loss": r["loss"],
    }
    print(json.dumps(out))
    if world > 1:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
