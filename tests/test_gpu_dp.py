"""Data-parallel training step on ONE GPU: G emulated ranks each run
kge_train_grad on their share of the batch, the buffers are summed the way the all-reduce
would, kge_train_apply updates — and the result must equal the oracle's step on the WHOLE
batch (same Philox negatives as a one-GPU step, same losses 1e-4, same weights)."""
import numpy as np
import pytest
import torch

from helpers import make_pair, tables_of, rand_triples, assert_weights_close
from oracle import ref_ops as O, ref_step as RS
from KGE._b200 import dp

pytestmark = pytest.mark.gpu
N, R = 300, 11

CASES = [
    ("TransE", O.LpDistance(1), O.PairwiseHingeLoss(1.0), 4, "h+t", 50, True),
    ("DistMult", None, O.BinaryCrossEntropyLoss(), 6, "h+t", 40, True),
    ("RotatE", O.LpDistance(1), O.SelfAdversarialNegativeSamplingLoss(3.0, 1.0), 8, "t", 24, True),
    ("TransH", O.LpDistancePow(2), O.PairwiseHingeLoss(1.0), 4, "h", 32, True),     # dense regulariser
    ("TransD", O.LpDistancePow(2), O.PairwiseLogisticLoss(), 2, "h+t", 20, True),
    ("UM", O.LpDistancePow(2), O.SquareErrorLoss(), 3, "t", 16, False),
]


@pytest.mark.parametrize("G", [1, 2, 3])
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_emulated_ranks_match_full_batch_oracle(case, G):
    model, score, loss, k, side, d, constraint = case
    rng = np.random.default_rng(5)
    ref, W, eng = make_pair(model, N, R, d, score=score, loss=loss, k=k, side=side, constraint=constraint)
    opt = RS.TFAdam()
    B, seed = 50, 21
    bufs = [eng.grad_buffer() for _ in range(G)]
    for step in range(3):
        X = rand_triples(rng, B, N, R)
        ids, k_eff = RS.sample_negatives(seed, step, X, k, side, N)
        l_ref, _ = RS.run_single_batch(ref, W, opt, X, ids, k_eff, side, True)
        negs = []
        for g in range(G):
            lo, hi = dp.split_even(B, G, g)
            negs.append(eng.train_grad(bufs[g], X[lo:hi], seed, step, B, lo, g, G, want_negatives=True))
        assert np.array_equal(torch.cat(negs).cpu().numpy(), ids), "ranks must draw the one-GPU negatives"
        for g in range(1, G):                         # what the all-reduce does
            bufs[0].flat += bufs[g].flat
        l_gpu = eng.train_apply(bufs[0], hi - lo)
        assert abs(float(l_gpu.item()) - l_ref) <= 1e-4 * max(1.0, abs(l_ref)), (step, float(l_gpu.item()), l_ref)
        got = tables_of(eng, model)
        for kk, v in W.data.items():
            assert_weights_close(got[kk], v, 1e-3, step + 1, (step, kk))


def test_one_rank_equals_fused_step_bitwise():
    """world = 1: grad + apply is the fused kge_train_step split in two (same record sums, same
    update arithmetic) — weights must be bit-identical."""
    rng = np.random.default_rng(9)
    _, _, a = make_pair("DistMult", N, R, 40, loss=O.BinaryCrossEntropyLoss(), k=6)
    _, _, b = make_pair("DistMult", N, R, 40, loss=O.BinaryCrossEntropyLoss(), k=6)
    gb = b.grad_buffer()
    for step in range(3):
        X = rand_triples(rng, 64, N, R)
        la = a.train_step(X, 3, step)
        b.train_grad(gb, X, 3, step, 64, 0)
        lb = b.train_apply(gb, 64)
        assert float(la.item()) == float(lb.item())
    for s in a.tab:
        assert torch.equal(a.tab[s].storage, b.tab[s].storage)


def test_rescal_with_dense_regulariser_is_refused():
    _, _, eng = make_pair("RESCAL", N, R, 12, loss=O.SquareErrorLoss(), constraint=True)
    with pytest.raises(RuntimeError, match="single-GPU"):
        eng.train_grad(eng.grad_buffer(), rand_triples(np.random.default_rng(0), 8, N, R), 1, 0, 8, 0)


MATRIX_CASES = [
    ("TransR", O.LpDistancePow(2), O.PairwiseHingeLoss(1.0), 4, "h+t", 24, 16, True),
    ("TransR", O.LpDistance(1), O.SelfAdversarialNegativeSamplingLoss(3.0, 1.0), 5, "t", 20, 28, True),
    ("SE", O.LpDistance(1), O.PairwiseLogisticLoss(), 3, "h+t", 20, None, True),
    ("RESCAL", None, O.SquareErrorLoss(), 4, "h", 16, None, False),
]


@pytest.mark.parametrize("G", [1, 2, 3])
@pytest.mark.parametrize("case", MATRIX_CASES, ids=[c[0] + str(i) for i, c in enumerate(MATRIX_CASES)])
def test_matrix_models_emulated_ranks_match_full_batch_oracle(case, G):
    """SE / TransR / RESCAL data parallel: every rank's k_update_mat writes its share of each relation matrix's
    gradient into the flat buffer (MODE 2), the buffers are summed as the all-reduce would, k_mat_apply steps the
    matrices — equal to the oracle's step on the whole batch (same negatives, losses 1e-4, weights at the parity
    tolerance); G = 1 against the fused step as well."""
    model, score, loss, k, side, d, dr, constraint = case
    rng = np.random.default_rng(5)
    ref, W, eng = make_pair(model, N, R, d, dr, score=score, loss=loss, k=k, side=side, constraint=constraint)
    opt = RS.TFAdam()
    B, seed = 48, 21
    bufs = [eng.grad_buffer() for _ in range(G)]
    for step in range(3):
        X = rand_triples(rng, B, N, R)
        ids, k_eff = RS.sample_negatives(seed, step, X, k, side, N)
        l_ref, _ = RS.run_single_batch(ref, W, opt, X, ids, k_eff, side, True)
        negs = []
        for g in range(G):
            lo, hi = dp.split_even(B, G, g)
            negs.append(eng.train_grad(bufs[g], X[lo:hi], seed, step, B, lo, g, G, want_negatives=True))
        assert np.array_equal(torch.cat(negs).cpu().numpy(), ids), "ranks must draw the one-GPU negatives"
        for g in range(1, G):
            bufs[0].flat += bufs[g].flat
        l_gpu = eng.train_apply(bufs[0], hi - lo)
        assert abs(float(l_gpu.item()) - l_ref) <= 1e-4 * max(1.0, abs(l_ref)), (step, float(l_gpu.item()), l_ref)
        got = tables_of(eng, model)
        for kk, v in W.data.items():
            assert_weights_close(got[kk], v, 1e-3, step + 1, (step, kk))


@pytest.mark.parametrize("G", [2, 3])
@pytest.mark.parametrize("case", [CASES[0], CASES[1], CASES[3]], ids=["TransE", "DistMult", "TransH"])
def test_peer_staging_equals_allreduce(case, G):
    """Fused exchange (kge_grads.peer): every emulated rank's gradient kernels store into all
    staging buffers, the apply half sums the slots in rank order.  Against the all-reduce
    emulation (sum of the per-rank buffers in rank order) the weights must be bit-identical —
    and every rank's staging must hold the same bits."""
    model, score, loss, k, side, d, constraint = case
    rng = np.random.default_rng(15)
    _, _, a = make_pair(model, N, R, d, score=score, loss=loss, k=k, side=side, constraint=constraint)
    _, _, b = make_pair(model, N, R, d, score=score, loss=loss, k=k, side=side, constraint=constraint)
    layout = {s: (pt.rows, pt.ld) for s, pt in b.tab.items()}
    probe = dp.PeerStaging(layout, G, 0)
    stage = [torch.full((probe.numel(),), 7.0, device=b.device) for _ in range(G)]   # one "rank memory" each
    ranks = [dp.PeerStaging(layout, G, g, base_ptrs=[t.data_ptr() for t in stage]) for g in range(G)]
    bufs = [a.grad_buffer() for _ in range(G)]
    B, seed = 48, 8
    for step in range(4):
        X = rand_triples(rng, B, N, R)
        for g in range(G):
            lo, hi = dp.split_even(B, G, g)
            a.train_grad(bufs[g], X[lo:hi], seed, step, B, lo, g, G)
            b.train_grad(ranks[g].push(step), X[lo:hi], seed, step, B, lo, g, G)
        acc = torch.zeros_like(bufs[0].flat)
        for g in range(G):
            acc += bufs[g].flat
        bufs[0].flat.copy_(acc)
        la = a.train_apply(bufs[0], hi - lo)
        par = step & 1
        half = G * probe.per
        for g in range(1, G):                     # all ranks received the same slots
            assert torch.equal(stage[0][par * half:(par + 1) * half], stage[g][par * half:(par + 1) * half])
        lb = b.train_apply(ranks[G - 1].slots(step), hi - lo)
        assert float(la.item()) == float(lb.item())
        for s in a.tab:
            assert torch.equal(a.tab[s].storage, b.tab[s].storage), (step, s)


@pytest.mark.parametrize("G", [2, 3, 8])
@pytest.mark.parametrize("case", [CASES[0], CASES[1], CASES[4]], ids=["TransE", "DistMult", "TransD"])
def test_owner_partitioned_exchange_equals_allreduce(case, G):
    """kge_grads.owner_rows: row sums go to the row's owner only, every emulated owner applies its rows
    and stores them into 'every replica' (one shared replica here).  Against the all-reduce emulation
    the weights — and the shadow table — must be bit-identical; m / v too (each row updated once)."""
    model, score, loss, k, side, d, constraint = case
    rng = np.random.default_rng(15)
    _, _, a = make_pair(model, N, R, d, score=score, loss=loss, k=k, side=side, constraint=constraint)
    _, _, b = make_pair(model, N, R, d, score=score, loss=loss, k=k, side=side, constraint=constraint)
    layout = {s: (pt.rows, pt.ld) for s, pt in b.tab.items()}
    ptabs = dp.PeerTables.local(layout, [s for s, pt in b.tab.items() if pt.wn is not None], G, b.device)
    b.adopt_peer_tables(ptabs)
    probe = dp.PeerStaging(layout, G, 0)
    stage = [torch.full((probe.numel(),), 7.0, device=b.device) for _ in range(G)]
    ranks = [dp.PeerStaging(layout, G, g, base_ptrs=[t.data_ptr() for t in stage], ptabs=ptabs) for g in range(G)]
    bufs = [a.grad_buffer() for _ in range(G)]
    B, seed = 48, 8
    for step in range(4):
        X = rand_triples(rng, B, N, R)
        if step == 2:                             # a validation batch in between (drops the shadow table)
            assert float(a.train_step(X, seed, 77, is_train=False).item()) == float(b.train_step(X, seed, 77, is_train=False).item())
        for g in range(G):
            lo, hi = dp.split_even(B, G, g)
            a.train_grad(bufs[g], X[lo:hi], seed, step, B, lo, g, G)
            b.train_grad(ranks[g].push(step), X[lo:hi], seed, step, B, lo, g, G)
        acc = torch.zeros_like(bufs[0].flat)
        for g in range(G):
            acc += bufs[g].flat
        bufs[0].flat.copy_(acc)
        la = a.train_apply(bufs[0], hi - lo)
        t0, valid0 = b.adam_t, {s: pt.wn_valid for s, pt in b.tab.items()}
        for g in range(G):                        # every owner applies its rows (each from the pre-step state)
            b.adam_t = t0
            for s, pt in b.tab.items():
                pt.wn_valid = valid0[s]
            lb = b.train_apply(ranks[g].slots(step), hi - lo)
        assert float(la.item()) == float(lb.item())
        for s in a.tab:
            assert torch.equal(a.tab[s].storage, b.tab[s].storage), (step, s)
            assert torch.equal(a.tab[s].m, b.tab[s].m) and torch.equal(a.tab[s].v, b.tab[s].v), (step, s)
        if a.tab[0].wn is not None:
            assert a.tab[0].wn_valid and b.tab[0].wn_valid
            assert torch.equal(a.tab[0].wn, b.tab[0].wn), step
