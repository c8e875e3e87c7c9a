"""Run under torchrun on >= 2 GPUs:
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/shard_check.py [--bench]
Real NVLink run of the sharded-entity-table step (kge_shard_*, KGE/_b200/shard.py) that
tests/test_gpu_shard.py emulates on one device:
  * losses / weights equal a one-GPU run over the same global batches (tolerances of the parity tests),
  * the gathered entity table is bitwise identical on every rank and between two runs,
  * sharded filtered ranking on the gathered table equals the one-GPU ranks exactly.
--bench: C5 timing (N = 5 M, d = 128, TransE L1 hinge 'h+t' k = 32, global B = 8192 x world), one JSON line."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "knowledge-graph-embedding_b200"), os.path.join(ROOT, "tests")]


def build(cfg, W_ent, W_rel, N, world, rank, dev, B_local, k_eff):
    from KGE._b200 import shard as SH
    S = SH.shard_rows(N, world)
    lds = {s: (v.shape[1] + 3) // 4 * 4 for s, v in W_ent.items()}
    mem = SH.ShardMemory.symmetric(S, lds, (2 + k_eff) * B_local, dev)
    lo, hi = rank * S, min(N, (rank + 1) * S)
    return SH.ShardedEngine(cfg, {s: v[lo:hi] for s, v in W_ent.items()}, W_rel, rank, world, mem, B_local, device=dev)


def check(rank, world, dev):
    from KGE._b200 import dp, engine as E, spec
    from oracle import ref_ops as O
    N, R, d, k = 3000, 17, 64, 8
    B = 96 * world
    g = torch.Generator().manual_seed(4)
    W = {0: (torch.rand((N, d), generator=g) - 0.5) * 0.6, 2: (torch.rand((R, d), generator=g) - 0.5) * 0.6}
    cfg = spec.make_cfg("TransE", N, R, d, score_fn=O.LpDistance(1), loss_fn=O.PairwiseHingeLoss(1.0), negative_ratio=k,
                        corrupt_side="h+t", constraint=True)
    ok = True
    tables = []
    for rep in range(2):
        rng = np.random.default_rng(1)
        single = E.Engine(cfg, W, device=dev)
        sh = build(cfg, {0: W[0]}, {2: W[2]}, N, world, rank, dev, B // world + 1, k)
        for step in range(6):
            X = np.stack([rng.integers(0, N, B), rng.integers(0, R, B), rng.integers(0, N, B)], 1)
            X[:, 2] = np.where(rng.random(B) < 0.4, X[:, 2] % 40, X[:, 2])
            lo, hi = dp.batch_shard(B)
            lm = sh.train_step(X[lo:hi], 11, step, B, lo)
            ls = single.train_step(X, 11, step)
            ok &= abs(float(lm.item()) - float(ls.item())) <= 1e-5 * max(1.0, abs(float(ls.item())))
        full = sh.gather_entity_table(0)
        ref = single.tab[0].view
        ok &= float((full - ref).abs().max()) <= 6 * 1e-3 * 1.01
        ok &= float(((full - ref).abs() > 1e-4 * float(ref.abs().max())).double().mean()) < 0.01
        gathered = [torch.empty_like(full) for _ in range(world)]
        dist.all_gather(gathered, full.contiguous())
        ok &= all(torch.equal(gathered[0], t) for t in gathered)
        ok &= bool(torch.equal(sh.eng.tab[2].view, single.tab[2].view)) or \
            float((sh.eng.tab[2].view - single.tab[2].view).abs().max()) <= 6 * 1e-3 * 1.01
        tables.append(full.clone())
        # sharded ranking on the gathered table == one-GPU ranks
        ev = E.Engine(cfg, {0: full, 2: sh.eng.tab[2].view}, device=dev)
        Q = np.stack([rng.integers(0, N, 200), rng.integers(0, R, 200), rng.integers(0, N, 200)], 1)
        lo_e, hi_e = dp.entity_shard(N)
        cnt = ev.rank_counts(Q, 1, ent_lo=lo_e, ent_hi=hi_e)
        dp.allreduce_counts(cnt)
        ok &= bool(torch.equal(cnt, ev.rank_counts(Q, 1)))
        del sh
    ok &= bool(torch.equal(tables[0], tables[1]))
    flag = torch.tensor([int(ok)], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("shard_check:", "PASS" if int(flag.item()) == 1 else "FAIL", f"(world {world})", flush=True)
    return int(flag.item()) == 1


def bench(rank, world, dev, steps=30, warm=4, N=5_000_000, R=1000, d=128, k=32, B_local=8192):
    from KGE._b200 import spec, shard as SH
    from oracle import ref_ops as O
    S = SH.shard_rows(N, world)
    lo, hi = rank * S, min(N, (rank + 1) * S)
    g = torch.Generator().manual_seed(100 + rank)
    lim = 6.0 / np.sqrt(d)
    W_ent = {0: (torch.rand((hi - lo, d), generator=g) * 2 - 1) * lim}
    W_rel = {2: (torch.rand((R, d), generator=torch.Generator().manual_seed(5)) * 2 - 1) * lim}
    out = []
    for constraint in (False, True):
        cfg = spec.make_cfg("TransE", N, R, d, score_fn=O.LpDistance(1), loss_fn=O.PairwiseHingeLoss(1.0), negative_ratio=k,
                            corrupt_side="h+t", constraint=constraint)
        lds = {0: d}
        mem = SH.ShardMemory.symmetric(S, lds, (2 + k) * B_local, dev)
        sh = SH.ShardedEngine(cfg, W_ent, W_rel, rank, world, mem, B_local, device=dev)
        rng = np.random.default_rng(77 + rank)
        wrel = 1.0 / np.arange(1, R + 1); wrel /= wrel.sum()
        batches = [torch.from_numpy(np.stack([rng.integers(0, N, B_local), rng.choice(R, size=B_local, p=wrel),
                                              rng.integers(0, N, B_local)], 1).astype(np.int64)).to(dev) for _ in range(8)]
        loss = torch.zeros(1, device=dev)
        B_global = B_local * world
        for i in range(warm):
            sh.train_step(batches[i % 8], 3, i, B_global, rank * B_local, loss_out=loss)
        torch.cuda.synchronize(); dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(steps):
            sh.train_step(batches[(warm + i) % 8], 3, warm + i, B_global, rank * B_local, loss_out=loss)
        b.record(); torch.cuda.synchronize()
        ms = torch.tensor([a.elapsed_time(b) / steps], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        # per-phase (host-synchronised, so only indicative)
        ph = {}
        for name in ("lookup", "pre", "grad", "allreduce", "apply"):
            ph[name] = 0.0
        sh.eng.profile_begin(32)
        for i in range(4):
            X = batches[i % 8]
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
            ev[0].record(); sh.lookup(X, 3, 100 + i, B_global, rank * B_local); sh.mem.barrier()
            ev[1].record(); sh.pre(); sh.mem.barrier()
            ev[2].record(); sh.grad()
            ev[3].record(); sh.gbuf.allreduce()
            ev[4].record(); sh.apply(loss)
            ev[5].record(); torch.cuda.synchronize()
            for j, name in enumerate(("lookup", "pre", "grad", "allreduce", "apply")):
                ph[name] += ev[j].elapsed_time(ev[j + 1]) / 4
        lib_ph, n_calls = sh.eng.profile_end()
        ph["lib_step_kernel"] = lib_ph["step"] / 4
        ph["lib_push_and_rel_update"] = lib_ph["update"] / 4 - ph["apply"]      # phase 4 = grad's push + apply's update
        ph["lib_sort_chain"] = lib_ph["sort"] / 4
        if rank == 0:
            ms = float(ms.item())
            alg = B_global * (7 * 4 * ((2 + k) * d + d) + 24)
            print(json.dumps({"config": f"C5 sharded entity table N={N} d={d} TransE L1 hinge h+t k={k} constraint={constraint}",
                              "n_gpus": world, "B_global": B_global, "B_local": B_local, "ms_per_step": round(ms, 4),
                              "train_triples_per_s": round(B_global * (1 + k) / (ms * 1e-3)),
                              "algorithmic_GBps_per_gpu": round(alg / world / (ms * 1e-3) / 1e9, 1),
                              "phase_ms_rank0": {kk: round(v, 4) for kk, v in ph.items()}, "final_loss": float(loss.item())}), flush=True)
        del sh, mem
        torch.cuda.empty_cache()


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ok = check(rank, world, dev)
    if "--bench" in sys.argv and ok:
        bench(rank, world, dev)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
