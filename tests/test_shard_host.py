"""Host logic of the sharded-entity-table step (KGE/_b200/shard.py) that needs no GPU: shard sizes,
the layout of the peer-visible buffer, the ctypes mirror of kge_shard."""
import ctypes as C

import pytest


def test_shard_rows_cover_the_table_and_stay_aligned():
    from KGE._b200 import shard as SH
    for n, w in [(5_000_000, 8), (14_951, 2), (1000, 3), (64, 8), (1, 1)]:
        S = SH.shard_rows(n, w)
        assert S % 64 == 0 and S * w >= n and (S - 64) * w < n + 64 * w
    assert SH.shard_rows(5_000_000, 8) == 625_024


def test_shard_memory_layout_sections_do_not_overlap():
    from KGE._b200 import shard as SH
    m = SH.ShardMemory(S=1024, lds={0: 128, 1: 132}, world_size=4, cap=300)
    secs = [(m.off_w[0], 1024 * 128), (m.off_w[1], 1024 * 132), (m.off_ids, 4 * 300), (m.off_cnt, 4),
            (m.off_val[0], 4 * 300 * 128), (m.off_val[1], 4 * 300 * 132)]
    secs.sort()
    for (a, n), (b, _) in zip(secs, secs[1:]):
        assert a + n <= b
    assert all(a % 64 == 0 for a, _ in secs)                      # 256-byte aligned sections (16-byte rows need it)
    assert secs[-1][0] + secs[-1][1] <= m.numel
    m.base_ptrs = [1 << 20, 1 << 30]
    assert m.addr(1, m.off_ids) == (1 << 30) + 4 * m.off_ids


def test_kge_shard_struct_matches_header():
    from KGE._b200 import lib
    assert C.sizeof(lib.Shard) == 4 + 4 + 8 + 8 + 8 * 2 * 8 + 8 * 8 + 8 * 8 + 8 * 2 * 8 + 4 * 8
    assert lib.Shard.w_peer.offset == 24 and lib.Shard.stamp.offset == 24 + 128 + 64 + 64 + 128


def test_sharded_engine_needs_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from KGE._b200 import shard as SH
    with pytest.raises(RuntimeError):
        m = SH.ShardMemory(64, {0: 8}, 1, 16)
        SH.ShardedEngine({"model": 2, "n_ent": 4, "n_rel": 2, "d_ent": 8, "d_rel": 8}, {}, {}, 0, 1, m, 4)
