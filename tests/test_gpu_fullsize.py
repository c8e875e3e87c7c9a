"""Parity at the FULL sizes of BASELINE.json's configurations (the small-shape tests cover the variants):
the CUDA step against the oracle on one or two whole batches, negatives against the Philox spec, and
size-independent properties (bitwise reproducibility, unit norms, filtered <= raw, tensor-core path ==
generic path, row-lazy == dense)."""
import os

import numpy as np
import pytest
import torch

from helpers import make_pair, tables_of, rand_triples, assert_weights_close, RankBand, oracle_rank_check
from oracle import ref_ops as O, ref_step as RS

pytestmark = pytest.mark.gpu


def _zipf(rng, n, N, R):
    def z(count, size):
        w = 1.0 / np.arange(1, count + 1)
        return rng.permutation(count)[rng.choice(count, size=size, p=w / w.sum())]
    return np.stack([z(N, n), z(R, n), z(N, n)], 1).astype(np.int64)


def test_c2_full_batch_matches_oracle_and_is_reproducible():
    N, R, d, B, k = 14951, 1345, 200, 1024, 64                    # BASELINE.json configs[1]
    rng = np.random.default_rng(1)
    batches = [_zipf(rng, B, N, R) for _ in range(3)]
    ref, W, eng = make_pair("DistMult", N, R, d, loss=O.BinaryCrossEntropyLoss(), k=k, side="h+t", constraint=True)
    _, _, eng2 = make_pair("DistMult", N, R, d, loss=O.BinaryCrossEntropyLoss(), k=k, side="h+t", constraint=True)
    opt = RS.TFAdam()
    for step, X in enumerate(batches):
        l_gpu, neg = eng.train_step(X, 12345, step, want_negatives=True)
        l2 = eng2.train_step(X, 12345, step)
        assert float(l_gpu.item()) == float(l2.item())                # bitwise reproducible
        if step < 2:                                                   # the oracle needs ~0.1 s per step here
            ids, k_eff = RS.sample_negatives(12345, step, X, k, "h+t", N)
            assert np.array_equal(neg.cpu().numpy(), ids), "negatives differ from the Philox spec"
            l_ref, _ = RS.run_single_batch(ref, W, opt, X, ids, k_eff, "h+t", True)
            assert abs(float(l_gpu.item()) - l_ref) <= 1e-4 * max(1.0, abs(l_ref)), (step, float(l_gpu.item()), l_ref)
    got = tables_of(eng, "DistMult")
    # the oracle stopped after two steps: compare the tables of a third engine run for two steps
    _, _, eng3 = make_pair("DistMult", N, R, d, loss=O.BinaryCrossEntropyLoss(), k=k, side="h+t", constraint=True)
    for step, X in enumerate(batches[:2]):
        eng3.train_step(X, 12345, step)
    got2 = tables_of(eng3, "DistMult")
    for kk, v in W.data.items():
        assert_weights_close(got2[kk], v, 1e-3, 2, kk)
    for s in eng.tab:
        assert torch.equal(eng.tab[s].storage, eng2.tab[s].storage)
    assert eng.tab[0].wn_valid                                         # shadow rows: unit norm (constraint.py:26-31)
    nrm = eng.tab[0].wn[:, :d].norm(dim=1)
    assert float((nrm - 1).abs().max()) < 1e-6
    assert all(torch.isfinite(v).all() for v in got.values())


def test_c2_full_ranking_properties():
    from KGE._b200.filter import build_filter
    N, R, d = 14951, 1345, 200
    rng = np.random.default_rng(2)
    ref, W, eng = make_pair("DistMult", N, R, d, loss=O.BinaryCrossEntropyLoss(), k=64, side="h+t", constraint=True)
    train = _zipf(rng, 200_000, N, R)
    Q = train[rng.integers(0, len(train), 512)]
    for side, sname in ((0, "h"), (1, "t")):
        off, ids = build_filter(train, Q, sname, N, eng.device)
        raw = eng.rank_counts(Q, side).cpu().numpy()
        flt = eng.rank_counts(Q, side, off, ids).cpu().numpy()
        assert (flt <= raw).all() and (flt >= 0).all() and raw.max() < N
        # entity shards sum to the whole (sharded evaluation), at an uneven split
        parts = sum(eng.rank_counts(Q, side, off, ids, ent_lo=lo, ent_hi=hi).cpu().numpy()
                    for lo, hi in ((0, 5000), (5000, 5001), (5001, N)))
        assert np.array_equal(parts, flt)
        os.environ["KGE_RANK_GENERIC"] = "1"                         # tensor-core path == generic fp32 path
        try:
            gen = eng.rank_counts(Q, side, off, ids).cpu().numpy()
        finally:
            os.environ.pop("KGE_RANK_GENERIC")
        assert np.array_equal(gen, flt)


def _train_a_bit(eng, ref_W, model, triples, B, steps, seed=99):
    """A few hundred GPU steps so the ranked scores are those of a (briefly) trained model, then the weights go
    back to the oracle's table dict."""
    rng = np.random.default_rng(seed)
    for step in range(steps):
        eng.train_step(triples[rng.integers(0, len(triples), B)], seed, step)
    for kk, v in tables_of(eng, model).items():
        ref_W.data[kk] = v.clone()


def _oracle_ranks_both_sides(ref, W, eng, queries, positive_X, n_ent, where):
    """>= 512 oracle-checked FILTERED queries per side (north_star: exact except documented 1e-5 ties)."""
    from KGE._b200.filter import build_filter
    out = {}
    for side, sname in ((0, "h"), (1, "t")):
        off, ids = build_filter(positive_X, queries, sname, n_ent, eng.device)
        flt = eng.rank_counts(queries, side, off, ids).cpu().numpy()
        band = RankBand()
        for q, x in enumerate(queries):
            oracle_rank_check(band, ref, W, x, sname, positive_X, int(flt[q]) + 1, (where, sname, q))
        print(where, sname, band.summary())
        assert band.used <= len(queries) // 50, band.summary()          # the band is the exception (<= 2 %)
        out[sname] = flt + 1
    return out


def test_c2_tensor_core_ranks_match_oracle_512_per_side():
    """configs[1] at full size on the tcgen05 path: N = 14 951, d = 200, 512 filtered queries per side
    against the oracle's get_rank (BaseModel.py:620-654), after 300 training steps."""
    N, R, d = 14951, 1345, 200
    rng = np.random.default_rng(22)
    ref, W, eng = make_pair("DistMult", N, R, d, loss=O.BinaryCrossEntropyLoss(), k=64, side="h+t", constraint=True)
    train = _zipf(rng, 200_000, N, R)
    _train_a_bit(eng, W, "DistMult", train, 1024, 300)
    Q = train[rng.integers(0, len(train), 512)].copy()
    Q[::5] = _zipf(rng, len(Q[::5]), N, R)                            # some queries outside positive_X
    _oracle_ranks_both_sides(ref, W, eng, Q, train, N, "C2 synthetic")


def _bundled(name):
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "data_bundled.npz"))
    return g[name].astype(np.int64)


def test_fb15k_valid_filtered_ranks_match_oracle():
    """Bundled data/fb15k/valid_indexed/valid.csv (50 000 triples, SURVEY.md 8d C2 rank parity; the train split is
    a missing blob, so the model trains on and filters with the valid split): DistMult d = 200, tcgen05 path."""
    N, R, d = 14951, 1345, 200
    valid = _bundled("fb15k_valid")
    ref, W, eng = make_pair("DistMult", N, R, d, loss=O.BinaryCrossEntropyLoss(), k=64, side="h+t", constraint=True)
    _train_a_bit(eng, W, "DistMult", valid, 1024, 300)
    rng = np.random.default_rng(23)
    Q = valid[rng.choice(len(valid), 512, replace=False)]
    _oracle_ranks_both_sides(ref, W, eng, Q, valid, N, "fb15k valid")


def test_fb15k237_valid_filtered_ranks_match_oracle():
    """configs[0] on the bundled FB15k-237: TransE L1 d = 50 trained on train_indexed, filtered ranks of 512 valid
    triples per side with positive_X = train U valid (BaseModel.py:646-654), CUDA-core tile path."""
    N, R, d = 14505, 237, 50
    train, valid = _bundled("fb15k237_train"), _bundled("fb15k237_valid")
    ref, W, eng = make_pair("TransE", N, R, d, score=O.LpDistance(1), loss=O.PairwiseHingeLoss(1.0), k=1, side="t")
    _train_a_bit(eng, W, "TransE", train, 4096, 300)
    rng = np.random.default_rng(24)
    Q = valid[rng.choice(len(valid), 512, replace=False)]
    ranks = _oracle_ranks_both_sides(ref, W, eng, Q, np.concatenate([train, valid]), N, "fb15k-237 valid")
    assert float(np.mean(1.0 / np.concatenate([ranks["h"], ranks["t"]]))) > 0.01      # it did learn something


def test_c1_full_batch_matches_oracle():
    N, R, d, B = 14505, 237, 50, 4096                                 # configs[0], headline batch
    rng = np.random.default_rng(3)
    ref, W, eng = make_pair("TransE", N, R, d, score=O.LpDistance(1), loss=O.PairwiseHingeLoss(1.0), k=1, side="t")
    opt = RS.TFAdam()
    for step in range(2):
        X = rand_triples(rng, B, N, R)
        ids, k_eff = RS.sample_negatives(7, step, X, 1, "t", N)
        l_ref, _ = RS.run_single_batch(ref, W, opt, X, ids, k_eff, "t", True)
        l_gpu, neg = eng.train_step(X, 7, step, want_negatives=True)
        assert np.array_equal(neg.cpu().numpy(), ids)
        assert abs(float(l_gpu.item()) - l_ref) <= 1e-4 * max(1.0, abs(l_ref))
    got = tables_of(eng, "TransE")
    for kk, v in W.data.items():
        assert_weights_close(got[kk], v, 1e-3, 2, kk)


def test_c5_shape_row_lazy_equals_dense_at_scale():
    """C5's shape scaled to 400 k entities (the dense twin must fit the test's time): B = 8192, k' = 32."""
    N, R, d, B = 400_000, 1000, 128, 8192
    outs = []
    for lazy in (False, True):
        rng = np.random.default_rng(4)
        _, _, eng = make_pair("TransE", N, R, d, score=O.LpDistance(1), loss=O.PairwiseHingeLoss(1.0), k=32, side="h+t",
                              constraint=True, lazy=lazy)
        ls = [float(eng.train_step(rand_triples(rng, B, N, R), 12345, step).item()) for step in range(5)]
        eng.flush()
        outs.append((ls, eng.tab[0].storage.clone(), eng.tab[0].v.clone()))
        del eng
    assert outs[0][0] == outs[1][0]
    assert torch.equal(outs[0][1], outs[1][1]) and torch.equal(outs[0][2], outs[1][2])


def test_c3_rotate_full_rows_and_negatives_match_oracle():
    """configs[2]: RotatE d = 500 (1000-float rows), self-adversarial, 256 negatives; 128 positives keep the
    oracle's autograd graph small, the row width / negative count / table shape are the full ones."""
    N, R, d, B, k = 14505, 237, 500, 128, 256
    rng = np.random.default_rng(5)
    loss = O.SelfAdversarialNegativeSamplingLoss(3.0, 1.0)
    ref, W, eng = make_pair("RotatE", N, R, d, score=O.LpDistance(1), loss=loss, k=k, side="h+t", constraint=True)
    opt = RS.TFAdam()
    for step in range(2):
        X = rand_triples(rng, B, N, R)
        ids, k_eff = RS.sample_negatives(5, step, X, k, "h+t", N)
        l_ref, _ = RS.run_single_batch(ref, W, opt, X, ids, k_eff, "h+t", True)
        l_gpu, neg = eng.train_step(X, 5, step, want_negatives=True)
        assert np.array_equal(neg.cpu().numpy(), ids)
        assert abs(float(l_gpu.item()) - l_ref) <= 1e-4 * max(1.0, abs(l_ref)), (step, float(l_gpu.item()), l_ref)
    got = tables_of(eng, "RotatE")
    for kk, v in W.data.items():
        assert_weights_close(got[kk], v, 1e-3, 2, kk)


@pytest.mark.parametrize("model", ["TransR", "TransD"])
def test_c4_typed_full_batch_matches_oracle(model):
    """configs[3]: d_e = d_r = 100, LpDistancePow(2), hinge, typed sampling over 32 types, 'h+t' k = 10, B = 1024."""
    from oracle import philox
    N, R, d, B, k = 14505, 237, 100, 1024, 10
    rng = np.random.default_rng(6)
    types = rng.integers(0, 32, N)
    pools = philox.TypedPools(ind2type=[int(t) for t in types])
    ref, W, eng = make_pair(model, N, R, d, d, score=O.LpDistancePow(2), loss=O.PairwiseHingeLoss(1.0), k=k, side="h+t",
                            constraint=True)
    eng.set_typed_pools(pools.ent_type, pools.offsets, pools.members, pools.pos)
    opt = RS.TFAdam()
    for step in range(2):
        X = rand_triples(rng, B, N, R)
        ids, k_eff = RS.sample_negatives(8, step, X, k, "h+t", N, typed=pools)
        l_ref, _ = RS.run_single_batch(ref, W, opt, X, ids, k_eff, "h+t", True)
        l_gpu, neg = eng.train_step(X, 8, step, want_negatives=True)
        assert np.array_equal(neg.cpu().numpy(), ids), "typed negatives differ from the spec"
        assert (types[ids.reshape(B, k_eff)[:, 0::2]] == types[X[:, 0]][:, None]).all()      # type preserved
        assert abs(float(l_gpu.item()) - l_ref) <= 1e-4 * max(1.0, abs(l_ref)), (step, float(l_gpu.item()), l_ref)
    got = tables_of(eng, model)
    for kk, v in W.data.items():
        assert_weights_close(got[kk], v, 1e-3, 2, kk)


def test_fb15k237_gpu_trained_vs_oracle_trained_metrics():
    """north_star "matching MRR/Hits@10": the same seeded epoch on the bundled FB15k-237 through the GPU step graph
    and through the CPU oracle; the two sets of weights then rank the same validation queries with their own code."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts"))
    import accuracy_parity
    r = accuracy_parity.run(epochs=1, n_queries=200, log=lambda *_: None)
    print({kk: r[kk] for kk in ("gpu", "cpu_oracle", "epoch_loss_max_rel_diff", "ranks_identical_fraction", "max_abs_weight_diff")})
    assert r["epoch_loss_max_rel_diff"] < 1e-4
    assert r["abs_diff"]["mrr"] < 5e-3 and r["abs_diff"]["hits10"] < 1e-2, r["abs_diff"]
    assert r["ranks_identical_fraction"] > 0.9
