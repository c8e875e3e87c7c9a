"""Standalone plug-in API (KGE.score / loss / constraint / metrics / data_utils): the
invariants of the reference's tests/test_score.py, test_loss.py, test_constraint.py,
test_metrics.py, test_data_utils.py, plus equality with the golden vectors produced by the
reference's own source.  CPU-only: these callables are descriptors on the hot path and plain
torch/numpy helpers when called directly."""
import math
import os

import numpy as np
import pytest
import torch

from KGE import score, loss, constraint, metrics, data_utils

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def close(a, b, tol=1e-5):
    np.testing.assert_allclose(np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64), rtol=tol, atol=tol)


def test_scores_match_reference_source():
    g = np.load(os.path.join(GOLD, "ops.npz"))
    x, y = g["x"], g["y"]
    cx, cy = torch.complex(torch.as_tensor(g["x"]), torch.as_tensor(g["xi"])), \
        torch.complex(torch.as_tensor(g["y"]), torch.as_tensor(g["yi"]))
    fns = {"lp1": score.LpDistance(1), "lp2": score.LpDistance(2), "lp3": score.LpDistance(3),
           "lpinf": score.LpDistance(np.inf), "lppow2": score.LpDistancePow(2), "lppow1": score.LpDistancePow(1),
           "dot": score.Dot()}
    for name, fn in fns.items():
        out = fn(x, y)
        assert len(out) == 10 and bool(torch.isfinite(out).all())
        if name != "dot":
            assert bool((out <= 0).all())                      # tests/test_score.py
            close(fn(cx, cy), g["cscore_" + name])
        close(out, g["score_" + name])
    assert score.SquaredLpDistance is score.LpDistancePow


def test_losses_match_reference_source():
    g = np.load(os.path.join(GOLD, "ops.npz"))
    fns = {"hinge": loss.PairwiseHingeLoss(margin=1.0), "hinge3": loss.PairwiseHingeLoss(margin=3.0),
           "logistic": loss.PairwiseLogisticLoss(), "bce": loss.BinaryCrossEntropyLoss(),
           "selfadv": loss.SelfAdversarialNegativeSamplingLoss(margin=3.0, temperature=1.0),
           "selfadv2": loss.SelfAdversarialNegativeSamplingLoss(margin=1.5, temperature=0.5),
           "square": loss.SquareErrorLoss()}
    for name, fn in fns.items():
        out = fn(g["pos"], g["neg"])
        assert out.dim() == 0 and bool(torch.isfinite(out)) and float(out) >= 0          # tests/test_loss.py
        close(out, g["loss_" + name])


def test_constraints_match_reference_source():
    g = np.load(os.path.join(GOLD, "ops.npz"))
    X = g["X"]
    for p, tag in [(1, "1"), (2, "2"), (np.inf, "inf")]:
        n = constraint.normalized_embeddings(X, p=p, value=1.0, axis=-1)
        c = constraint.clip_constraint(X, p=p, value=1.0, axis=-1)
        close(n, g["norm_p" + tag], 1e-6)
        close(c, g["clip_p" + tag], 1e-6)
        nn = torch.linalg.vector_norm(n, ord=p, dim=-1)
        assert n.shape == X.shape and bool(((nn - 1).abs() <= 1e-6).all())           # tests/test_constraint.py
        assert bool((torch.linalg.vector_norm(c, ord=p, dim=-1) - 1 <= 1e-6).all())
    close(constraint.soft_constraint(X, p=2, value=1.0, axis=-1), g["soft_p2"])
    close(constraint.Lp_regularization(X, p=2, axis=-1), g["lpreg_p2"])
    close(constraint.Lp_regularization(X, p=3, axis=-1), g["lpreg_p3"])


def test_metrics_ranges_and_golden():
    rng = np.random.default_rng(0)
    ranks = rng.integers(1, 1001, 10)
    vals = {"mr": metrics.mean_rank(ranks), "mrr": metrics.mean_reciprocal_rank(ranks),
            "med": metrics.median_rank(ranks), "g": metrics.geometric_mean_rank(ranks),
            "h": metrics.harmonic_mean_rank(ranks), "std": metrics.std_rank(ranks), "hit": metrics.hits_at_k(ranks, 10)}
    assert all(isinstance(v, float) for v in vals.values())                          # tests/test_metrics.py
    assert 0 < vals["mrr"] <= 1 and vals["mr"] >= 1 and vals["med"] >= 1 and vals["g"] >= 1 and vals["h"] >= 1
    assert vals["std"] >= 0 and vals["hit"] <= 1
    g = np.load(os.path.join(GOLD, "case_transe.npz"))          # scipy gmean/hmean of the reference, same ranks
    r = g["rank_flt_t"]
    mine = [metrics.mean_rank(r), metrics.mean_reciprocal_rank(r), metrics.median_rank(r),
            metrics.geometric_mean_rank(r), metrics.harmonic_mean_rank(r), metrics.std_rank(r),
            metrics.hits_at_k(r, 1), metrics.hits_at_k(r, 3), metrics.hits_at_k(r, 10)]
    close(mine, g["eval_vals"], 1e-9)
    with pytest.raises(AssertionError):
        metrics.hits_at_k(r, 0)
    # evaluate()'s one-pass summary: the same nine values in the reference's key order (BaseModel.py:606-618)
    s = metrics.summary(r)
    assert list(s) == ["mean_rank", "mean_reciprocal_rank", "median_rank", "geometric_mean_rank", "harmonic_mean_rank",
                       "std_rank", "hit@1", "hit@3", "hit@10"] and list(s.values()) == mine
    for n in (1, 2, 3, 10, 11, 1000, 1001):                     # the partition median is np.median's value
        x = rng.integers(1, 50, n)
        assert metrics.median_rank(x) == float(np.median(x))
    assert all(np.isnan(v) for v in metrics.summary(np.zeros(0, dtype=np.int64)).values())


def test_index_and_split(tmp_path):
    from toy_kg import train_raw, train, metadata
    assert [metadata["ent2ind"][e] for e in metadata["ind2ent"]] == list(range(len(metadata["ind2ent"])))
    assert [metadata["rel2ind"][r] for r in metadata["ind2rel"]] == list(range(len(metadata["ind2rel"])))
    assert train.shape == train_raw.shape and train.max() < len(metadata["ind2ent"])
    # csv folder round trip (tests/test_data_utils.py)
    folder = tmp_path / "kg"
    folder.mkdir()
    np.savetxt(folder / "part0.csv", train_raw[:7], fmt="%s", delimiter=",")
    np.savetxt(folder / "part1.csv", train_raw[7:], fmt="%s", delimiter=",")
    md = data_utils.index_kg(str(folder))
    data_utils.convert_kg_to_index(str(folder), md["ent2ind"], md["rel2ind"])
    back = data_utils.load_triples(str(folder) + "_indexed")
    assert back.dtype == np.int32 and len(back) == len(train_raw) == data_utils.calculate_data_size(str(folder))
    assert [md["ind2ent"][i] for i in back[:, 0]] == list(train_raw[:, 0])
    big = np.random.default_rng(1).integers(0, 30, (400, 3))
    tr, te = data_utils.train_test_split_no_unseen(big, test_size=0.1, seed=3)
    assert len(te) == 40 and len(tr) == 360
    assert set(te[:, 0]) | set(te[:, 2]) <= set(tr[:, 0]) | set(tr[:, 2]) and set(te[:, 1]) <= set(tr[:, 1])


def test_batch_stream_semantics_cpu():
    """shuffle(buffer=n) -> repeat -> batch: exact batch size, batches straddle passes, each pass
    is a permutation; unshuffled data wraps around."""
    data = np.arange(30).reshape(10, 3)
    it = data_utils.DeviceBatchStream(data, 4, shuffle=True, seed=5, device="cpu")
    rows = torch.cat([next(it) for _ in range(5)])           # 20 rows = two passes
    assert rows.shape == (20, 3) and rows.dtype == torch.int64
    assert sorted(rows[:10, 0].tolist()) == sorted(data[:, 0].tolist())
    assert sorted(rows[10:, 0].tolist()) == sorted(data[:, 0].tolist())
    it2 = data_utils.DeviceBatchStream(data, 4, shuffle=True, seed=5, device="cpu")
    assert torch.equal(torch.cat([next(it2) for _ in range(5)]), rows)
    w = data_utils.DeviceBatchStream(data, 4, shuffle=False, device="cpu")
    got = torch.cat([next(w) for _ in range(3)])
    assert got[:, 0].tolist() == [0, 3, 6, 9, 12, 15, 18, 21, 24, 27, 0, 3]


def test_train_test_split_follows_the_reference_on_duplicate_draws():
    """/root/reference/KGE/data_utils.py:139-154 restated: an index drawn twice is tested and its entity / relation
    counters are decremented again before np.unique collapses it.  With the same seed the mirror must return the
    same split (the mirror used to skip repeated draws)."""
    import numpy as np
    from KGE import data_utils

    def reference_split(X, test_size, seed):
        if isinstance(test_size, float):
            test_size = int(len(X) * test_size)
        e, e_cnt = np.unique(np.append(X[:, 0], X[:, 2]), return_counts=True)
        r, r_cnt = np.unique(X[:, 1], return_counts=True)
        e_dict, r_dict = dict(zip(e, e_cnt)), dict(zip(r, r_cnt))
        test_id, train_id = np.array([], dtype=int), np.arange(len(X))
        loop_count, max_loop = 0, len(X) * 10
        rnd = np.random.RandomState(seed)
        while len(test_id) < test_size:
            i = rnd.choice(train_id)
            if e_dict[X[i, 0]] > 1 and r_dict[X[i, 1]] > 1 and e_dict[X[i, 2]] > 1:
                e_dict[X[i, 0]] -= 1
                r_dict[X[i, 1]] -= 1
                e_dict[X[i, 2]] -= 1
                test_id = np.unique(np.append(test_id, i))
            loop_count += 1
            if loop_count == max_loop:
                return None
        return X[np.setdiff1d(train_id, test_id)], X[test_id]

    rng = np.random.default_rng(0)
    for n, N, R, ts in [(400, 40, 5, 0.4), (300, 25, 4, 120), (1000, 60, 7, 0.3)]:
        X = np.stack([rng.integers(0, N, n), rng.integers(0, R, n), rng.integers(0, N, n)], 1)
        for seed in (1, 2, 3):
            want, got = reference_split(X, ts, seed), data_utils.train_test_split_no_unseen(X, ts, seed)
            assert (want is None) == (got is None)
            if want is not None:
                assert np.array_equal(want[0], got[0]) and np.array_equal(want[1], got[1])
