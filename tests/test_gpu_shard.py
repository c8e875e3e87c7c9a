"""Entity tables sharded over ranks (kge_shard_*), ranks emulated on one device:
the owner-applies protocol (ids routed to owners, peer gathers, gradient rows pushed to owners, rank-order
sum) against the single-GPU step on the whole batch."""
import numpy as np
import pytest
import torch

from helpers import make_pair, rand_triples, assert_weights_close
from oracle import ref_ops as O, ref_models as RM

pytestmark = pytest.mark.gpu

CASES = [
    ("TransE", O.LpDistance(1), O.PairwiseHingeLoss(1.0), True, 50, 4, "h+t"),
    ("DistMult", None, O.BinaryCrossEntropyLoss(), True, 72, 6, "h+t"),
    ("TransD", O.LpDistancePow(2), O.PairwiseHingeLoss(1.0), True, 40, 4, "h+t"),
    ("RotatE", O.LpDistance(1), O.SelfAdversarialNegativeSamplingLoss(3.0, 1.0), True, 65, 6, "t"),
    ("UM", O.LpDistance(2), O.PairwiseLogisticLoss(), True, 33, 3, "h"),
]
IDS = [c[0] for c in CASES]
N, R = 1000, 6


def _sharded(case, world, B_local_max, seed=7):
    from KGE._b200 import spec, shard as SH
    model, score, loss, con, d, k, side = case
    ref = RM.MODELS[model](N, R, d, None, score_fn=score, loss_fn=loss, constraint=con, constraint_weight=1.0)
    W = ref.init_weights(seed)
    cfg = spec.make_cfg(model, N, R, d, None, score_fn=score, loss_fn=loss, negative_ratio=k, corrupt_side=side,
                        constraint=con, rotate_limit=ref.limit)
    slots = spec.SLOTS[model]
    S = SH.shard_rows(N, world)
    ent = {slots[kk]: v for kk, v in W.data.items() if slots[kk] < 2}
    rel = {slots[kk]: v for kk, v in W.data.items() if slots[kk] >= 2}
    lds = {s: (int(np.prod(v.shape[1:])) + 3) // 4 * 4 for s, v in ent.items()}
    k_eff = 2 * (k // 2) if side == "h+t" else k
    mems = SH.ShardMemory.local(S, lds, world, (2 + k_eff) * B_local_max, "cuda:0")
    ranks = []
    for g in range(world):
        lo, hi = g * S, min(N, (g + 1) * S)
        ranks.append(SH.ShardedEngine(cfg, {s: v[lo:hi] for s, v in ent.items()}, rel, g, world, mems[g], B_local_max))
    return ranks


def _full_tables(ranks):
    out = {}
    for s in ranks[0].eng.tab:
        if s < 2:
            out[s] = torch.cat([r.eng.tab[s].view for r in ranks], 0).cpu()
        else:
            out[s] = ranks[0].eng.tab[s].view.cpu()
    return out


def _batches(rng, steps, B):
    out = []
    for _ in range(steps):
        X = rand_triples(rng, B, N, R)
        X[:, 2] = np.where(rng.random(B) < 0.4, X[:, 2] % 30, X[:, 2])       # rows shared between ranks
        out.append(X)
    return out


@pytest.mark.parametrize("case", CASES, ids=IDS)
def test_one_rank_shard_equals_lazy_step_bitwise(case):
    from KGE._b200 import shard as SH
    model, score, loss, con, d, k, side = case
    B = 64
    batches = _batches(np.random.default_rng(3), 6, B)
    _, _, eng = make_pair(model, N, R, d, score=score, loss=loss, k=k, side=side, constraint=con, lazy=True)
    ranks = _sharded(case, 1, B)
    for step, X in enumerate(batches):
        l_ref = eng.train_step(X, 5, step)
        l_sh = SH.emulate_step(ranks, [X], 5, step, B, [0])
        assert float(l_ref.item()) == float(l_sh.item()), step
    full = _full_tables(ranks)
    for s, pt in eng.tab.items():
        assert torch.equal(pt.view.cpu(), full[s]), f"slot {s}"
    # Adam slots of the shard as well
    ranks[0].eng.flush(); eng.flush()
    for s, pt in eng.tab.items():
        assert torch.equal(pt.m, ranks[0].eng.tab[s].m) and torch.equal(pt.v, ranks[0].eng.tab[s].v), f"slot {s} m/v"


@pytest.mark.parametrize("case", CASES, ids=IDS)
@pytest.mark.parametrize("world", [2, 3])
def test_sharded_ranks_match_single_gpu_step(case, world):
    from KGE._b200 import shard as SH, dp
    model, score, loss, con, d, k, side = case
    B, steps = 96, 8
    batches = _batches(np.random.default_rng(4), steps, B)
    _, _, eng = make_pair(model, N, R, d, score=score, loss=loss, k=k, side=side, constraint=con)
    runs = []
    for rep in range(2):
        ranks = _sharded(case, world, B)
        losses = []
        for step, X in enumerate(batches):
            parts = [dp.split_even(B, world, g) for g in range(world)]
            if rep == 0:
                l_ref, neg_ref = eng.train_step(X, 5, step, want_negatives=True)
            l_sh = SH.emulate_step(ranks, [X[lo:hi] for lo, hi in parts], 5, step, B, [lo for lo, _ in parts])
            losses.append(float(l_sh.item()))
            if rep == 0:
                lr = float(l_ref.item())
                assert abs(losses[-1] - lr) <= 1e-5 * max(1.0, abs(lr)), (step, losses[-1], lr)
        full = _full_tables(ranks)
        runs.append((losses, full))
        if rep == 0:
            for s, pt in eng.tab.items():
                assert_weights_close(full[s], pt.view.cpu(), 1e-3, steps, where=f"slot {s}")
    assert runs[0][0] == runs[1][0]                                   # bitwise reproducible
    for s in runs[0][1]:
        assert torch.equal(runs[0][1][s], runs[1][1][s]), f"slot {s} differs between two runs"


def test_shard_inbox_capacity_is_checked():
    from KGE._b200 import shard as SH
    ranks = _sharded(CASES[0], 2, 8)
    X = rand_triples(np.random.default_rng(0), 64, N, R)
    with pytest.raises(RuntimeError, match="cap"):
        ranks[0].lookup(X, 1, 0, 64, 0)


@pytest.mark.parametrize("case", [CASES[0], CASES[2]], ids=[IDS[0], IDS[2]])
def test_fused_push_equals_separate_push_bitwise(case):
    """The step kernel storing single-record rows straight into the owner's inbox (default) against
    k_shard_push delivering every row (KGE_OPT_NO_FUSED_PUSH)."""
    from KGE._b200 import shard as SH, dp, lib as L
    B, world = 96, 3
    batches = _batches(np.random.default_rng(8), 5, B)
    runs = []
    for nofuse in (1, 0):
        ranks = _sharded(case, world, B)
        ranks[0].eng.set_option(L.OPT_NO_FUSED_PUSH, nofuse)
        try:
            losses = []
            for step, X in enumerate(batches):
                parts = [dp.split_even(B, world, g) for g in range(world)]
                losses.append(float(SH.emulate_step(ranks, [X[lo:hi] for lo, hi in parts], 5, step, B, [lo for lo, _ in parts]).item()))
            runs.append((losses, _full_tables(ranks)))
        finally:
            ranks[0].eng.set_option(L.OPT_NO_FUSED_PUSH, 0)
    assert runs[0][0] == runs[1][0]
    for s in runs[0][1]:
        assert torch.equal(runs[0][1][s], runs[1][1][s]), f"slot {s}"


@pytest.mark.parametrize("opt", ["sgd", "adagrad"])
def test_one_rank_shard_equals_lazy_step_bitwise_sgd_adagrad(opt):
    """SGD / Adagrad on sharded tables: the owner's row update takes kge_cfg.optimizer."""
    from KGE._b200 import shard as SH
    case = CASES[0]
    model, score, loss, con, d, k, side = case
    B = 64
    batches = _batches(np.random.default_rng(3), 6, B)
    _, _, eng = make_pair(model, N, R, d, score=score, loss=loss, k=k, side=side, constraint=con, lazy=True)
    ranks = _sharded(case, 1, B)
    for e in [eng] + [r.eng for r in ranks]:
        e.cfg.optimizer, e.cfg.lr = (1 if opt == "sgd" else 2), 0.05
        if opt == "adagrad":
            e.acc_init = 0.1
    for step, X in enumerate(batches):
        l_ref = eng.train_step(X, 5, step)
        l_sh = SH.emulate_step(ranks, [X], 5, step, B, [0])
        assert float(l_ref.item()) == float(l_sh.item()), step
    eng.flush()
    for r in ranks:
        r.eng.flush()
    full = _full_tables(ranks)
    for s, pt in eng.tab.items():
        assert torch.equal(pt.view.cpu(), full[s]), s
