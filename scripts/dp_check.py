"""Run under torchrun on >= 2 GPUs: python -m torch.distributed.run --nproc-per-node 2 scripts/dp_check.py
Checks on real NCCL what tests/test_gpu_dp.py emulates on one device:
  * replicas stay bit-identical over data-parallel steps,
  * the weights equal a one-GPU run over the same global batches (tolerance of the parity tests),
  * entity-sharded evaluate() returns the one-GPU ranks exactly."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "knowledge-graph-embedding_b200"), os.path.join(ROOT, "tests")]


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from KGE._b200 import dp, engine as E, spec, filter as F
    from oracle import ref_ops as O
    N, R, d, B, k = 2000, 37, 64, 96 * world, 8
    rng = np.random.default_rng(1)
    g = torch.Generator().manual_seed(4)
    W = {"ent_emb": (torch.rand((N, d), generator=g) - 0.5) * 0.6, "rel_emb": (torch.rand((R, d), generator=g) - 0.5) * 0.6}
    cfg = spec.make_cfg("TransE", N, R, d, score_fn=O.LpDistance(1), loss_fn=O.PairwiseHingeLoss(1.0), negative_ratio=k,
                        corrupt_side="h+t", constraint=True)
    slots = spec.SLOTS["TransE"]
    multi = E.Engine(cfg, {slots[n]: v for n, v in W.items()}, device=dev)
    single = E.Engine(cfg, {slots[n]: v for n, v in W.items()}, device=dev)
    nccl = E.Engine(cfg, {slots[n]: v for n, v in W.items()}, device=dev)      # same steps over the NCCL all-reduce
    owner = E.Engine(cfg, {slots[n]: v for n, v in W.items()}, device=dev)     # owner-partitioned exchange
    ahead = E.Engine(cfg, {slots[n]: v for n, v in W.items()}, device=dev)     # owner exchange + look-ahead grouping
    graph = E.Engine(cfg, {slots[n]: v for n, v in W.items()}, device=dev)     # the same step as a captured graph
    graph.dp_w_private = True                                                    # ... with only wn sent to the replicas
    ok = True
    Xs = [np.stack([rng.integers(0, N, B), rng.integers(0, R, B), rng.integers(0, N, B)], 1) for _ in range(5)]
    lo, hi = dp.batch_shard(B)
    Xl = [torch.as_tensor(x[lo:hi]).to(dev) for x in Xs]
    os.environ["KGE_DP_EXCHANGE"] = "owner"
    G = graph.step_graph_dp(hi - lo, 11, B, lo)
    os.environ.pop("KGE_DP_EXCHANGE")
    ok &= G is not None
    for step in range(5):
        X = Xs[step]
        os.environ["KGE_DP_EXCHANGE"] = "owner"
        nxt = Xl[step + 1] if step + 1 < 5 and step != 2 else None
        la = ahead.train_step_dp(Xl[step], 11, step, B, lo, next_triples=nxt)
        lg = G.launch(Xl[step], step, next_triples=nxt)
        os.environ.pop("KGE_DP_EXCHANGE")
        os.environ["KGE_DP_EXCHANGE"] = "peer"
        lm = multi.train_step_dp(X[lo:hi], 11, step, B, lo)
        os.environ["KGE_DP_EXCHANGE"] = "owner"
        lo_ = owner.train_step_dp(X[lo:hi], 11, step, B, lo)
        ok &= float(lm.item()) == float(lo_.item())          # both sum the rank slots in rank order
        ok &= float(la.item()) == float(lo_.item()) and float(lg.item()) == float(lo_.item())   # look-ahead / graph: same bits
        os.environ["KGE_DP_EXCHANGE"] = "nccl"
        ln = nccl.train_step_dp(X[lo:hi], 11, step, B, lo)
        os.environ.pop("KGE_DP_EXCHANGE")
        if world == 2:       # a + b in either order: the two exchanges must agree bit for bit
            ok &= float(lm.item()) == float(ln.item())
        ls = single.train_step(X, 11, step)
        ok &= abs(float(lm.item()) - float(ls.item())) <= 1e-4 * max(1.0, abs(float(ls.item())))
    try:
        graph.tab[0].view
        ok = False                      # stale replicas must refuse to be read
    except RuntimeError:
        pass
    graph.dp_sync_weights()
    for s in multi.tab:
        mine = multi.tab[s].storage
        ref = single.tab[s].storage
        ok &= float((mine - ref).abs().max()) <= 5 * 1e-3 * 1.01
        ok &= float(((mine - ref).abs() > 1e-4 * float(ref.abs().max())).double().mean()) < 0.01
        gathered = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(gathered, mine)
        ok &= all(torch.equal(gathered[0], t) for t in gathered)            # replicas bit-identical
        if world == 2:
            ok &= bool(torch.equal(mine, nccl.tab[s].storage))
        ok &= bool(torch.equal(mine, owner.tab[s].storage))   # owner-partitioned == store-to-all, any world size
        ok &= bool(torch.equal(ahead.tab[s].storage, owner.tab[s].storage)) and bool(torch.equal(graph.tab[s].storage, owner.tab[s].storage))
        go = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(go, owner.tab[s].storage.contiguous())
        ok &= all(torch.equal(go[0], t) for t in go)
    if rank == 0:
        print("exchange kinds:", multi._dp_kind, nccl._dp_kind, owner._dp_kind, ahead._dp_kind, graph._dp_kind)
    # sharded evaluation: counts over this rank's entity slice, summed == one-GPU counts
    Q = np.stack([rng.integers(0, N, 300), rng.integers(0, R, 300), rng.integers(0, N, 300)], 1)
    posX = np.concatenate([Q, np.stack([rng.integers(0, N, 4000), rng.integers(0, R, 4000), rng.integers(0, N, 4000)], 1)])
    for side, sid in (("h", 0), ("t", 1)):
        off, ids = F.build_filter(torch.as_tensor(posX).to(dev), Q, side, N, dev)
        e_lo, e_hi = dp.entity_shard(N)
        part = dp.allreduce_counts(multi.rank_counts(Q, sid, off, ids, ent_lo=e_lo, ent_hi=e_hi))
        full = multi.rank_counts(Q, sid, off, ids)
        ok &= bool(torch.equal(part, full))
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("DP_CHECK", "PASS" if int(flag.item()) == 1 else "FAIL", "world", world)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
