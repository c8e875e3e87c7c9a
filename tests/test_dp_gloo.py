"""Host logic of the multi-GPU paths over gloo, world_size 2, on CPU:
shard arithmetic, the flat gradient buffer's single all-reduce, and the sharded-count
reduction of evaluation.  The kernels themselves need a GPU (tests/test_gpu_dp.py emulates the
ranks on one device); here the per-rank work is done with torch CPU ops."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from KGE._b200 import dp


def test_split_even_covers_range_without_overlap():
    for n in (0, 1, 7, 1024, 14951):
        for parts in (1, 2, 3, 8):
            bounds = [dp.split_even(n, parts, i) for i in range(parts)]
            assert bounds[0][0] == 0 and bounds[-1][1] == n
            assert all(bounds[i][1] == bounds[i + 1][0] for i in range(parts - 1))
            sizes = [hi - lo for lo, hi in bounds]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        dp.split_even(10, 2, 2)


def test_world_defaults_to_single_process():
    assert dp.world() == (0, 1)
    assert dp.entity_shard(100) == (0, 100)
    assert dp.batch_shard(64) == (0, 64)
    c = torch.arange(5, dtype=torch.int32)
    assert torch.equal(dp.allreduce_counts(c.clone()), c)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        res = {}
        # ---- evaluation: entity-sharded strict-greater counts, summed ------------------------
        g = torch.Generator().manual_seed(3)
        N, d, Q = 101, 8, 17
        E = torch.randn(N, d, generator=g)
        q = torch.randn(Q, d, generator=g)
        pos = (q * E[torch.arange(Q) % N]).sum(1)
        lo, hi = dp.entity_shard(N)
        res["shard"] = (lo, hi)
        cnt = ((q @ E[lo:hi].T) > pos[:, None]).sum(1).to(torch.int32)
        res["counts"] = dp.allreduce_counts(cnt).tolist()
        res["counts_full"] = ((q @ E.T) > pos[:, None]).sum(1).tolist()
        # ---- training: batch shards + the flat gradient buffer ------------------------------
        B = 33
        b_lo, b_hi = dp.batch_shard(B)
        res["batch"] = (b_lo, b_hi)
        gb = dp.GradBuffer({0: (N, 8), 2: (5, 12)}, "cpu")
        assert gb.view(0).shape == (N, 8) and gb.view(2).shape == (5, 12) and gb.scal.numel() == 8
        rows = torch.arange(b_lo, b_hi)
        gb.view(0).index_add_(0, rows % N, torch.ones(len(rows), 8) * (rank + 1))
        gb.view(2)[rank] = float(rank + 1)
        gb.scal[0] = float(b_hi - b_lo)               # "norm^2" share
        gb.scal[4] = 0.5 * (rank + 1)                 # loss share
        gb.allreduce()
        res["g0_sum"] = float(gb.view(0).sum())
        res["g2"] = gb.view(2)[:, 0].tolist()
        res["scal"] = gb.scal.tolist()
        res["flat_hash"] = float(gb.flat.double().mul(torch.arange(gb.flat.numel()).double() + 1).sum())
        out[rank] = res
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_two_ranks_over_gloo():
    world, port = 2, _free_port()
    with mp.Manager() as m:
        out = m.dict()
        mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
        r0, r1 = out[0], out[1]
    # shards tile the ranges
    assert r0["shard"] == (0, 51) and r1["shard"] == (51, 101)
    assert r0["batch"] == (0, 17) and r1["batch"] == (17, 33)
    # summed shard counts equal the single-process counts, identically on both ranks
    assert r0["counts"] == r0["counts_full"] == r1["counts"]
    # one all-reduce carries every table and the scalars; both replicas hold the same bits
    assert r0["flat_hash"] == r1["flat_hash"]
    assert r0["g0_sum"] == 17 * 8 * 1 + 16 * 8 * 2
    assert r0["g2"][:2] == [1.0, 2.0]
    assert r0["scal"][0] == 33.0 and r0["scal"][4] == 1.5


def test_peer_staging_addresses():
    """Host logic of the fused exchange: rank r pushes into slot r of every rank's staging, and
    reads slot j of its own; parities alternate between two disjoint halves."""
    from KGE._b200 import dp
    layout = {0: (10, 8), 2: (3, 8)}
    G = 3
    bases = [1 << 20, 2 << 20, 3 << 20]
    ranks = [dp.PeerStaging(layout, G, r, base_ptrs=bases) for r in range(G)]
    per = ranks[0].per
    assert per % 64 == 0 and per >= 10 * 8 + 3 * 8 + 8 and ranks[0].numel() == 2 * G * per + 64
    for par in (0, 1):
        for r in range(G):
            P, S = ranks[r].push(par), ranks[r].slots(par)
            assert P.n_peers == G and S.n_peers == G
            for j in range(G):
                # what r pushes to j is what j reads as slot r
                Sj = ranks[j].slots(par)
                for s in layout:
                    assert P.peer[j][s] == Sj.peer[r][s]
                assert P.peer_scal[j] == Sj.peer_scal[r]
                assert bases[r] <= S.peer[j][0] < bases[r] + 4 * ranks[r].numel()
    a0, a1 = ranks[1].slots(0), ranks[1].slots(1)
    assert a1.peer[0][0] - a0.peer[0][0] == 4 * G * per
    assert dp.exchange_kind(2) == "owner" and dp.exchange_kind(8) == "owner" and dp.exchange_kind(1) == "nccl"
    assert dp.exchange_kind(2, dense_regulariser=True) == "peer" and dp.exchange_kind(8, dense_regulariser=True) == "nccl"
    F = ranks[1].peer_flags()                       # kge_peer_barrier's flag words: after both parities of every rank
    assert F.world == G and F.rank == 1
    assert [F.flags[j] for j in range(G)] == [bases[j] + 4 * 2 * G * per for j in range(G)]


def test_owner_partitioned_exchange_host_logic():
    """kge_grads.owner_rows / w_peer / wn_peer as PeerStaging fills them, and the PeerTables layout."""
    from KGE._b200 import dp
    layout = {0: (10, 8), 2: (3, 8)}
    G = 4
    pt = dp.PeerTables(layout, wn_slots=[0])
    assert set(pt.off) == {("w", 0), ("wn", 0), ("w", 2)}
    offs = sorted(v[0] for v in pt.off.values())
    assert offs == [0, 128, 256] and pt.numel == 320              # 256-byte aligned sections
    pt.base_ptrs = [j << 24 for j in range(1, G + 1)]
    st = [dp.PeerStaging(layout, G, r, base_ptrs=[j << 20 for j in range(1, G + 1)], ptabs=pt) for r in range(G)]
    assert st[0].owner_rows == {0: 3, 2: 1}                         # ceil(rows / world)
    for r in range(G):
        for Gs in (st[r].push(0), st[r].slots(1)):
            assert Gs.owner_rank == r and list(Gs.owner_rows) == [3, 0, 1, 0]
            for j in range(G):
                assert Gs.w_peer[j][0] == pt.base_ptrs[j] + 4 * pt.off[("w", 0)][0]
                assert Gs.w_peer[j][2] == pt.base_ptrs[j] + 4 * pt.off[("w", 2)][0]
                assert Gs.wn_peer[j] == pt.base_ptrs[j] + 4 * pt.off[("wn", 0)][0]
    # every row has exactly one owner and the owners tile the table
    for s, (rows, _) in layout.items():
        owners = [r // st[0].owner_rows[s] for r in range(rows)]
        assert max(owners) < G and owners == sorted(owners)
    plain = dp.PeerStaging(layout, G, 0, base_ptrs=[1 << 20] * G)   # store-to-all form: owner fields stay zero
    assert list(plain.push(0).owner_rows) == [0, 0, 0, 0]
