"""torchrun --nproc-per-node N scripts/dp_phases.py — where the time of a data-parallel C2 step goes
(owner-partitioned exchange): CUDA events around the gradient half, the barriers and the apply half."""
import json, os, sys
import numpy as np, torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "knowledge-graph-embedding_b200"), os.path.join(ROOT, "tests")]
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
from KGE._b200 import engine as E, spec
from oracle import ref_ops as O
N, R, d, B, k = 14951, 1345, 200, 1024, 64
g = torch.Generator().manual_seed(4)
W = {0: (torch.rand((N, d), generator=g) - 0.5) * 0.2, 2: (torch.rand((R, d), generator=g) - 0.5) * 0.2}
cfg = spec.make_cfg("DistMult", N, R, d, loss_fn=O.BinaryCrossEntropyLoss(), negative_ratio=k, corrupt_side="h+t", constraint=True)
eng = E.Engine(cfg, W, device=dev)
rng = np.random.default_rng(rank)
X = [torch.from_numpy(np.stack([rng.integers(0, N, B), rng.integers(0, R, B), rng.integers(0, N, B)], 1)).to(dev) for _ in range(8)]
loss = torch.zeros(1, device=dev)
for i in range(10):
    eng.train_step_dp(X[i % 8], 3, i, B * world, rank * B, loss_out=loss)
names = ["grad", "barrier1", "apply", "barrier2"]
acc = dict.fromkeys(names, 0.0)
reps = 50
for i in range(reps):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
    par = eng._dp_steps & 1
    eng._dp_steps += 1
    ev[0].record(); eng.train_grad(eng._peer.push(par), X[i % 8], 3, 10 + i, B * world, rank * B, rank, world)
    ev[1].record(); eng._peer.barrier()
    ev[2].record(); eng.train_apply(eng._peer.slots(par), B, loss)
    ev[3].record(); eng._peer.barrier()
    ev[4].record(); torch.cuda.synchronize()
    for j, n in enumerate(names):
        acc[n] += ev[j].elapsed_time(ev[j + 1]) * 1e3 / reps
# whole steps enqueued back to back: GPU time per step and host enqueue time per step, both paths
res = {}
for one_call in ("1", "0"):
    os.environ["KGE_DP_ONE_CALL"] = one_call
    for i in range(5):
        eng.train_step_dp(X[i % 8], 3, 100 + i, B * world, rank * B, loss_out=loss)
    torch.cuda.synchronize(); dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    import time
    a.record(); t0 = time.perf_counter()
    for i in range(200):
        eng.train_step_dp(X[i % 8], 3, 200 + i, B * world, rank * B, loss_out=loss)
    host_us = (time.perf_counter() - t0) * 1e6 / 200
    b.record(); torch.cuda.synchronize()
    res["one_call" if one_call == "1" else "four_calls"] = {"gpu_us_per_step": round(a.elapsed_time(b) * 1e3 / 200, 1), "host_enqueue_us_per_step": round(host_us, 1)}
t = torch.tensor([acc[n] for n in names], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    print(json.dumps({"world": world, "exchange": eng._dp_kind, "us_max_over_ranks": dict(zip(names, [round(float(x), 1) for x in t])), "back_to_back_rank0": res}))
dist.barrier(); dist.destroy_process_group()
