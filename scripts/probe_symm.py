import os, sys, torch, torch.distributed as dist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
try:
    import torch.distributed._symmetric_memory as sm
    t = sm.empty(1 << 20, dtype=torch.float32, device=dev)
    h = sm.rendezvous(t, group=dist.group.WORLD)
    if rank == 0:
        print("symm ok; attrs:", [a for a in dir(h) if not a.startswith("_")])
        print("buffer_ptrs", [hex(p) for p in h.buffer_ptrs], "world", h.world_size, "rank", h.rank)
    t.fill_(rank + 1.0)
    h.barrier()
    peer = h.get_buffer((rank + 1) % world, (16,), torch.float32)
    v = float(peer[0].item())
    print("rank", rank, "reads peer value", v)
    # time barriers
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(100): h.barrier()
    b.record(); torch.cuda.synchronize()
    if rank == 0: print("barrier us", a.elapsed_time(b) * 10)
    # time a 13 MB all_reduce
    x = torch.zeros(3300000, device=dev)
    for _ in range(5): dist.all_reduce(x)
    torch.cuda.synchronize(); a.record()
    for _ in range(50): dist.all_reduce(x)
    b.record(); torch.cuda.synchronize()
    if rank == 0: print("nccl all_reduce 13MB us", a.elapsed_time(b) * 20)
except Exception as e:
    print("rank", rank, "symm FAILED:", repr(e)[:500])
dist.destroy_process_group()
