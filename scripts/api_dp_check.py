"""torchrun --nproc-per-node N scripts/api_dp_check.py — the KGE model API under torch.distributed:
train() (owner-partitioned exchange by default) leaves bit-identical replicas on every rank, evaluate()
with entity-sharded ranking returns the same metrics everywhere."""
import os, sys, tempfile
import numpy as np, torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "knowledge-graph-embedding_b200"), os.path.join(ROOT, "tests")]
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
from KGE.models.translating_based.TransE import TransE
from KGE.models.semantic_based.DistMult import DistMult
from KGE.models.translating_based.TransR import TransR
rng = np.random.default_rng(0)
N, R = 500, 9
X = np.stack([rng.integers(0, N, 4000), rng.integers(0, R, 4000), rng.integers(0, N, 4000)], 1)
train, val = X[:3600], X[3600:]
md = {"ind2ent": [f"e{i}" for i in range(N)], "ent2ind": {f"e{i}": i for i in range(N)},
      "ind2rel": [f"r{i}" for i in range(R)], "rel2ind": {f"r{i}": i for i in range(R)}}
ok = True
for cls, kw, with_val in ((TransE, {}, True), (DistMult, {}, True), (DistMult, {}, False), (TransR, {}, True)):
    # without validation batches the owners keep w to themselves during an epoch (kge_grads.w_private);
    # TransR: matrix gradients in the all-reduced flat buffer (exchange "nccl")
    ep = {"ent_embedding_size": 32, "rel_embedding_size": 24} if cls is TransR else {"embedding_size": 32}
    m = cls(embedding_params=ep, negative_ratio=4, corrupt_side="h+t", **kw)
    m.train(train, val if with_val else None, md, epochs=3, batch_size=256, seed=5, log_path=tempfile.mkdtemp(),
            early_stopping_rounds=None)
    ok &= bool(getattr(m._engine._peer, "w_private", False)) == (not with_val and m._engine._dp_kind == "owner")
    kind = m._engine._dp_kind
    for k, v in m.model_weights.items():
        t = v.tensor.contiguous()
        g = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(g, t)
        ok &= all(torch.equal(g[0], x) for x in g) and bool(torch.isfinite(t).all())
    ev = m.evaluate(val[:64], "t", positive_X=X)
    objs = [None] * world
    dist.all_gather_object(objs, {k: float(v) for k, v in ev.items()})
    ok &= all(o == objs[0] for o in objs)
    ok &= np.isfinite(m.train_loss_history[-1]) and m.train_loss_history[-1] < m.train_loss_history[0]
    if rank == 0:
        print(cls.__name__, "exchange", kind, "w_private", not with_val, "loss", [round(x, 4) for x in m.train_loss_history], "mrr", round(objs[0].get("mrr", objs[0].get("MRR", 0)), 4), flush=True)
flag = torch.tensor([int(ok)], device="cuda")
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("API_DP_CHECK", "PASS" if int(flag.item()) else "FAIL", "world", world, flush=True)
dist.barrier(); dist.destroy_process_group()
sys.exit(0 if int(flag.item()) else 1)
