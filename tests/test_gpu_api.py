"""The drop-in KGE Python API on the GPU: mirrors the reference's tests/test_model.py,
tests/test_ns_strategy.py and tests/test_integration.py (same invariants, same combination
matrix), plus golden parity of the API path itself (model classes -> C-ABI)."""
import itertools
import math
import os

import numpy as np
import pytest
import torch

from KGE import score, loss, ns_strategy, optimizers
from KGE.models.translating_based.UM import UM
from KGE.models.translating_based.SE import SE
from KGE.models.translating_based.TransE import TransE
from KGE.models.translating_based.TransH import TransH
from KGE.models.translating_based.TransR import TransR
from KGE.models.translating_based.TransD import TransD
from KGE.models.translating_based.RotatE import RotatE
from KGE.models.semantic_based.DistMult import DistMult
from KGE.models.semantic_based.RESCAL import RESCAL
from toy_kg import train, val, metadata
from test_golden_oracle import load_case, weights, close, TOL

pytestmark = pytest.mark.gpu

TRANSLATING = {"UM": UM, "SE": SE, "TransE": TransE, "TransH": TransH, "TransR": TransR, "TransD": TransD,
               "RotatE": RotatE}
SEMANTIC = {"DistMult": DistMult, "RESCAL": RESCAL}
ALL = dict(TRANSLATING, **SEMANTIC)


def emb(name, d=16):
    return {"ent_embedding_size": d, "rel_embedding_size": d} if name in ("TransR", "TransD") else {"embedding_size": d}


def typed_metadata():
    md = dict(metadata)
    n = len(md["ind2ent"])
    md["ind2type"] = ["A"] * (n // 2) + ["B"] * (n - n // 2)
    md["type2inds"] = {t: np.array([i for i, ti in enumerate(md["ind2type"]) if ti == t]) for t in ("A", "B")}
    return md


# ---- tests/test_model.py -----------------------------------------------------------------
class TestBaseModel:
    def setup_method(self):
        self.model = TransE(embedding_params={"embedding_size": 16}, negative_ratio=2, corrupt_side="h+t")
        self.model._model_weights_initial = None
        self.model.metadata = metadata
        self.model._init_embeddings(seed=None)

    def test_train(self, tmp_path):
        self.model.train(train_X=train, val_X=val, metadata=metadata, epochs=1, batch_size=2, log_path=str(tmp_path))
        assert len(self.model.train_loss_history) == 1 and math.isfinite(self.model.train_loss_history[0])

    def test_get_rank(self):
        x = val[:1]
        rank = self.model.get_rank(x=x, positive_X=None, corrupt_side="h")
        rank_filter = self.model.get_rank(x=x, positive_X=np.concatenate((train, val), axis=0), corrupt_side="h")
        assert isinstance(rank, np.int_) and isinstance(rank_filter, np.int_)
        assert rank >= 1 and rank_filter >= 1 and rank_filter <= rank

    def test_evaluate(self):
        raw = self.model.evaluate(eval_X=val, corrupt_side="h", positive_X=None)
        flt = self.model.evaluate(eval_X=val, corrupt_side="h", positive_X=np.concatenate((train, val), axis=0))
        for m in ["mean_rank", "median_rank", "geometric_mean_rank", "harmonic_mean_rank"]:
            assert flt[m] <= raw[m]
        for m in ["mean_reciprocal_rank", "hit@1", "hit@3", "hit@10"]:
            assert flt[m] >= raw[m]
        assert list(raw.keys()) == ["mean_rank", "mean_reciprocal_rank", "median_rank", "geometric_mean_rank",
                                    "harmonic_mean_rank", "std_rank", "hit@1", "hit@3", "hit@10"]


@pytest.mark.parametrize("name", list(ALL))
def test_init_and_score_hrt(name):
    kw = dict(embedding_params=emb(name), negative_ratio=2, corrupt_side="h+t")
    model = ALL[name](**kw)
    model._model_weights_initial = None
    model.metadata = metadata
    model._init_embeddings(seed=None)
    n_ent, n_rel = len(metadata["ind2ent"]), len(metadata["ind2rel"])
    assert set(model.model_weights) == set(model._tables(n_ent, n_rel))
    for k, (shape, _) in model._tables(n_ent, n_rel).items():
        assert model.model_weights[k].shape == tuple(shape) and model.model_weights[k].numpy().dtype == np.float32
    s = model.score_hrt(h=val[:, 0], r=val[:, 1], t=val[:, 2])
    assert len(s) == len(val) and bool(torch.isfinite(s).all())
    assert len(model.score_hrt(h=None, r=val[0, 1], t=val[0, 2])) == n_ent
    assert len(model.score_hrt(h=val[0, 0], r=val[0, 1], t=None)) == n_ent
    with pytest.raises(AssertionError):
        model.score_hrt(h=None, r=val[0, 1], t=None)              # the reference's guard never fires


# ---- tests/test_ns_strategy.py ---------------------------------------------------------
def test_uniform_strategy_shape_dtype():
    n_ent = len(metadata["ind2ent"])
    strat = ns_strategy.UniformStrategy(sample_pool=np.arange(n_ent))
    for side in ("h", "t"):
        out = strat(torch.as_tensor(val), negative_ratio=2, side=side)
        assert tuple(out.shape) == (len(val) * 2,) and out.dtype == torch.int64
        assert int(out.min()) >= 0 and int(out.max()) < n_ent
    out32 = strat(torch.as_tensor(val.astype(np.int32)), negative_ratio=2, side="h")
    assert out32.dtype == torch.int32
    sub = ns_strategy.UniformStrategy(sample_pool=np.array([3, 5, 7]))
    assert set(sub(val, negative_ratio=50, side="t").cpu().tolist()) <= {3, 5, 7}


def test_typed_strategy_preserves_type():
    md = typed_metadata()
    strat = ns_strategy.TypedStrategy(pool=None, metadata=md)
    types = np.asarray(md["ind2type"])
    for side, col in (("h", 0), ("t", 2)):
        out = strat(torch.as_tensor(val), negative_ratio=2, side=side)
        assert tuple(out.shape) == (len(val) * 2,) and out.dtype == torch.int64
        ref = np.repeat(val[:, col], 2)
        got = out.cpu().numpy()
        assert (types[got] == types[ref]).all() and (got != ref).all()
    lonely = dict(md, ind2type=["Z"] + ["A"] * (len(types) - 1))
    lonely.pop("type2inds")
    with pytest.raises(ValueError):
        ns_strategy.TypedStrategy(pool=None, metadata=lonely)(val, negative_ratio=1, side="h")


# ---- tests/test_integration.py: every model x score x loss x sampler ----------------------
ALL_SCORES = [score.LpDistance(p=2), score.LpDistancePow(p=2), score.Dot()]
ALL_LOSSES = [loss.PairwiseHingeLoss(margin=1.0), loss.PairwiseLogisticLoss(), loss.BinaryCrossEntropyLoss(),
              loss.SquareErrorLoss(), loss.SelfAdversarialNegativeSamplingLoss(margin=1.0, temperature=1.0)]


@pytest.mark.parametrize("name", list(ALL))
def test_integration_matrix(name, tmp_path):
    md = typed_metadata()
    samplers = [ns_strategy.UniformStrategy(np.arange(len(md["ind2ent"]))), ns_strategy.TypedStrategy(pool=None, metadata=md)]
    scores = [None] if name in SEMANTIC else (ALL_SCORES[:2] if name == "RotatE" else ALL_SCORES)
    for sc, ls, smp in itertools.product(scores, ALL_LOSSES, samplers):
        kw = dict(embedding_params=emb(name), negative_ratio=2, corrupt_side="h+t", loss_fn=ls, ns_strategy=smp)
        if sc is not None:
            kw["score_fn"] = sc
        model = ALL[name](**kw)
        model.train(train_X=train, val_X=val, metadata=md, epochs=1, batch_size=4, early_stopping_rounds=None,
                    restore_best_weight=False, optimizer=optimizers.Adam(), seed=12345, log_path=str(tmp_path),
                    log_projector=True)
        res = model.evaluate(eval_X=val, corrupt_side="h")
        assert math.isfinite(model.train_loss_history[-1]) and res["mean_rank"] >= 1


def test_sampler_class_and_early_stopping(tmp_path):
    md = typed_metadata()
    for cls in (ns_strategy.UniformStrategy, ns_strategy.TypedStrategy):
        m = TransE(embedding_params={"embedding_size": 8}, negative_ratio=2, corrupt_side="t", ns_strategy=cls)
        m.train(train, val, md, epochs=6, batch_size=4, early_stopping_rounds=2, seed=1, log_path=str(tmp_path))
        assert 1 <= len(m.val_loss_history) <= 6 and m.best_step is not None
    m2 = TransE(embedding_params={"embedding_size": 8}, negative_ratio=1, corrupt_side="h+t")
    with pytest.raises(RuntimeError):          # 'h+t' with negative_ratio=1 -> zero negatives
        m2.train(train, None, md, epochs=1, batch_size=4, log_path=str(tmp_path))


def test_train_without_validation_and_restore(tmp_path):
    m = DistMult(embedding_params={"embedding_size": 8}, negative_ratio=4, corrupt_side="h+t")
    m.train(train, None, metadata, epochs=2, batch_size=4, seed=3, log_path=str(tmp_path))   # reference crashes (C.1)
    saved = {k: v.numpy().copy() for k, v in m.model_weights.items()}
    m.train(train, None, metadata, epochs=1, batch_size=4, seed=4, log_path=str(tmp_path))
    m.restore_model_weights(saved)                                                           # reference: TypeError (C.12)
    for k, v in saved.items():
        assert np.array_equal(m.model_weights[k].numpy(), v)
    same = DistMult(embedding_params={"embedding_size": 8}, negative_ratio=4, corrupt_side="h+t")
    same.train(train, None, metadata, epochs=2, batch_size=4, seed=3, log_path=str(tmp_path))
    same2 = {k: v.numpy() for k, v in same.model_weights.items()}
    m.train(train, None, metadata, epochs=2, batch_size=4, seed=3, log_path=str(tmp_path))
    for k in same2:                                    # seeded runs are bitwise reproducible end to end
        assert np.array_equal(m.model_weights[k].numpy(), same2[k])


# ---- the API path against the reference's own train()/evaluate() run (golden) -------------
def test_api_step_sequence_matches_reference_golden():
    """model class -> spec -> C-ABI: same losses as the reference's __run_single_batch sequence."""
    for name in ("transe", "distmult", "transr", "rotate", "rescal", "transh"):
        g, spec = load_case(name)
        cls = ALL[spec["model"]]
        sc = {"l1": score.LpDistance(1), "l2": score.LpDistance(2), "l2pow": score.LpDistancePow(2), "dot": score.Dot(),
              "None": None}[spec["score"]]
        ls = {"hinge": loss.PairwiseHingeLoss(1.0), "bce": loss.BinaryCrossEntropyLoss(), "square": loss.SquareErrorLoss(),
              "selfadv": loss.SelfAdversarialNegativeSamplingLoss(3.0, 1.0)}[spec["loss"]]
        kw = dict(embedding_params=emb(spec["model"], spec["d"]), negative_ratio=spec["k"], corrupt_side=spec["side"], loss_fn=ls)
        if spec["model"] in ("TransR", "TransD"):
            kw["embedding_params"] = {"ent_embedding_size": spec["d"], "rel_embedding_size": spec["dr"]}
        if sc is not None:
            kw["score_fn"] = sc
        if spec["model"] != "RotatE":
            kw["constraint"] = spec["constraint"]
        if spec["model"] in ("TransH", "DistMult", "RESCAL"):
            kw["constraint_weight"] = spec["cw"]
        m = cls(**kw)
        m.metadata = {"ind2ent": list(range(40)), "ind2rel": list(range(5))}
        m._model_weights_initial = {k: v.numpy() for k, v in weights(g, "w0_").items()}
        # w0 is the reference's post-init state: skip our init-time constraint by building the engine directly
        m._build_engine({k: torch.as_tensor(v) for k, v in m._model_weights_initial.items()})
        m._setup_sampler()
        for step in range(len(g["step_X"])):
            l = m._engine.train_step(g["step_X"][step], 4242, step, is_train=bool(g["step_is_train"][step]))
            want = float(g["step_loss"][step])
            assert abs(float(l.item()) - want) <= TOL * max(1.0, abs(want)), (name, step)
        sc_all = m.score_hrt(h=None, r=g["score_X"][0][1], t=g["score_X"][0][2])
        assert len(sc_all) == 40


@pytest.mark.gpu
def test_train_step_host_equals_device_step():
    """kge_train_step_host (pinned host batch in, host loss out) is the device step plus two
    copies: same weights bit for bit, same loss."""
    import numpy as np
    import torch
    from helpers import make_pair, rand_triples
    from oracle import ref_ops as O
    rng = np.random.default_rng(4)
    _, _, a = make_pair("DistMult", 300, 11, 40, loss=O.BinaryCrossEntropyLoss(), k=6)
    _, _, b = make_pair("DistMult", 300, 11, 40, loss=O.BinaryCrossEntropyLoss(), k=6)
    host_loss = torch.zeros(1, dtype=torch.float32).pin_memory()
    for step in range(3):
        X = rand_triples(rng, 64, 300, 11)
        la = a.train_step(X, 9, step)
        b.train_step_host(torch.from_numpy(X).pin_memory(), 9, step, host_loss)
        torch.cuda.synchronize()
        assert float(la.item()) == float(host_loss.item())
    for s in a.tab:
        assert torch.equal(a.tab[s].storage, b.tab[s].storage)


# ---- sharded entity tables through the model API (ranks emulated on one device) --------------
def _random_kg(n_ent=600, n_rel=7, n_train=4000, n_val=300, seed=3):
    rng = np.random.default_rng(seed)
    tri = lambda n: np.stack([rng.integers(0, n_ent, n), rng.integers(0, n_rel, n), rng.integers(0, n_ent, n)], 1)   # noqa: E731
    md = {"ind2ent": ["e%d" % i for i in range(n_ent)], "ind2rel": ["r%d" % i for i in range(n_rel)]}
    return tri(n_train).astype(np.int64), tri(n_val).astype(np.int64), md


@pytest.mark.parametrize("name,ranks", [("TransE", 2), ("DistMult", 3), ("TransD", 2)])
def test_train_with_sharded_tables_on_emulated_ranks(name, ranks, tmp_path, monkeypatch):
    """KGEModel.train picks the sharded step (kge_shard_*) when the entity tables are sharded; losses, weights and
    filtered metrics equal the replicated single-GPU run of the same model / seed."""
    tr, va, md = _random_kg()
    out = {}
    for mode in ("single", "sharded"):
        monkeypatch.setenv("KGE_B200_SHARD", "1" if mode == "sharded" else "0")
        monkeypatch.setenv("KGE_B200_SHARD_EMULATE", str(ranks) if mode == "sharded" else "0")
        m = ALL[name](embedding_params=emb(name, 24), negative_ratio=4, corrupt_side="h+t")
        m.train(train_X=tr, val_X=va, metadata=md, epochs=2, batch_size=250, seed=11, log_path=str(tmp_path / mode))
        assert (m._sharded is not None) == (mode == "sharded")
        res = m.evaluate(va[:100], "t", positive_X=np.concatenate([tr, va]))
        out[mode] = (m.train_loss_history, m.val_loss_history, {k: v.numpy() for k, v in m.model_weights.items()}, res,
                     m.score_hrt(va[:50, 0], va[:50, 1], va[:50, 2]).cpu().numpy())
    a, b = out["single"], out["sharded"]
    np.testing.assert_allclose(b[0], a[0], rtol=1e-5)
    np.testing.assert_allclose(b[1], a[1], rtol=1e-4)
    n_steps = 2 * 16
    for k in a[2]:
        diff = np.abs(a[2][k] - b[2][k])
        assert diff.max() <= n_steps * 1e-3 * 1.01 and (diff > 1e-4 * np.abs(a[2][k]).max()).mean() < 0.01, (k, diff.max())
    np.testing.assert_allclose(b[4], a[4], rtol=2e-3, atol=2e-3)
    assert abs(a[3]["mean_reciprocal_rank"] - b[3]["mean_reciprocal_rank"]) < 0.02
    # the checkpoint holds every rank's optimizer shard
    assert (tmp_path / "sharded" / "ckpt.pt").exists() and (tmp_path / "sharded" / "ckpt.opt.rank0.pt").exists()


def test_shard_and_lazy_are_chosen_for_every_optimizer(monkeypatch, tmp_path):
    """Row-lazy / sharded tables cover Adam, SGD and Adagrad (round 1 was Adam-only and the automatic choice had to
    step around the others): the choice depends on model and table size alone, and a forced row-lazy run with SGD
    trains through the model API."""
    from KGE._b200 import lib as L
    from KGE import optimizers
    from toy_kg import train, val, metadata
    m = TransE(embedding_params={"embedding_size": 128}, negative_ratio=2, corrupt_side="h+t")
    m.batch_size = 64
    big = {"ent_emb": torch.empty((2_000_000, 128), device="meta"), "rel_emb": torch.empty((10, 128), device="meta")}
    for opt in (L.OPTIM_ADAM, L.OPTIM_SGD, L.OPTIM_ADAGRAD):
        cfg = {"constraint": 1, "optimizer": opt}
        assert m._lazy_rows(cfg, big) is True
        monkeypatch.setenv("KGE_B200_SHARD_EMULATE", "2")
        assert bool(m._shard_world(cfg, big)) is True
        monkeypatch.delenv("KGE_B200_SHARD_EMULATE")
    monkeypatch.setenv("KGE_B200_LAZY_ROWS", "1")
    for opt in (optimizers.SGD(0.05), optimizers.Adagrad(0.05)):
        m2 = TransE(embedding_params={"embedding_size": 8}, negative_ratio=2, corrupt_side="h+t")
        m2.train(train, val, metadata, epochs=2, batch_size=4, optimizer=opt, seed=1, log_path=str(tmp_path))
        assert m2._engine.lazy and np.isfinite(m2.train_loss_history[-1])


# ---- checkpoints with optimizer state + resume -------------------------------------------------
@pytest.mark.parametrize("lazy", ["0", "1"])
def test_resume_continues_bitwise(lazy, tmp_path, monkeypatch):
    """4 epochs in one go == 2 epochs, a fresh model object, resume=True for the remaining 2: weights, Adam slots
    and losses bit for bit (the reference's checkpoints hold the weights only, BaseModel.py:247-253)."""
    monkeypatch.setenv("KGE_B200_LAZY_ROWS", lazy)
    tr, va, md = _random_kg(n_ent=300, n_train=1500, n_val=100)
    kw = dict(train_X=tr, val_X=va, metadata=md, batch_size=128, seed=5)
    full = TransE(embedding_params={"embedding_size": 16}, negative_ratio=4, corrupt_side="h+t")
    full.train(epochs=4, log_path=str(tmp_path / "full"), **kw)
    part = TransE(embedding_params={"embedding_size": 16}, negative_ratio=4, corrupt_side="h+t")
    part.train(epochs=2, log_path=str(tmp_path / "part"), **kw)
    again = TransE(embedding_params={"embedding_size": 16}, negative_ratio=4, corrupt_side="h+t")
    again.train(epochs=4, log_path=str(tmp_path / "part"), resume=True, **kw)
    assert again.train_loss_history == full.train_loss_history and again.val_loss_history == full.val_loss_history
    for k in full.model_weights:
        assert np.array_equal(full.model_weights[k].numpy(), again.model_weights[k].numpy()), k
    ef, ea = full._engine, again._engine
    ef.flush(); ea.flush()
    assert ef.adam_t == ea.adam_t
    for s in ef.tab:
        assert torch.equal(ef.tab[s].m, ea.tab[s].m) and torch.equal(ef.tab[s].v, ea.tab[s].v), s


def test_background_checkpoint_equals_the_synchronous_one(tmp_path, monkeypatch):
    """One GPU: the checkpoint is a pinned snapshot in stream order written by a worker thread under the next epoch
    (BaseModel._CheckpointWriter).  Its files are complete when train() returns and hold exactly what the
    synchronous writer (KGE_B200_ASYNC_CKPT=0) saves — weights, Adam slots, counters."""
    tr, va, md = _random_kg(n_ent=300, n_train=1500, n_val=100)
    kw = dict(train_X=tr, val_X=None, metadata=md, batch_size=128, seed=9, epochs=3)
    out = {}
    for mode in ("1", "0"):
        monkeypatch.setenv("KGE_B200_ASYNC_CKPT", mode)
        m = TransE(embedding_params={"embedding_size": 16}, negative_ratio=4, corrupt_side="h+t")
        m.train(log_path=str(tmp_path / mode), **kw)
        assert sorted(f for f in os.listdir(tmp_path / mode) if f.startswith("ckpt")) == ["ckpt.opt.rank0.pt", "ckpt.pt"]
        out[mode] = (torch.load(tmp_path / mode / "ckpt.pt"), torch.load(tmp_path / mode / "ckpt.opt.rank0.pt"), m)
    (ca, oa, ma), (cs, os_, ms) = out["1"], out["0"]
    assert ca["adam_t"] == cs["adam_t"] and ca["epoch"] == cs["epoch"] == 2 and ca["step"] == cs["step"]
    assert ca["train_loss_history"] == cs["train_loss_history"]
    for k in cs["weights"]:
        assert torch.equal(ca["weights"][k], cs["weights"][k]), k
        assert np.array_equal(ca["weights"][k].numpy(), ma.model_weights[k].numpy()), k        # the state after the last step
    for ea, es in zip(oa["engines"], os_["engines"]):
        assert ea["adam_t"] == es["adam_t"]
        for s2 in es["m"]:
            assert torch.equal(ea["m"][s2], es["m"][s2]) and torch.equal(ea["v"][s2], es["v"][s2]), s2


def test_device_histograms_match_numpy(tmp_path):
    """_Writers.histograms counts tf.summary.histogram's 30 equal-width buckets on the device; the counts are those
    of numpy on the same edges (values on an edge may fall either side: float32 edge arithmetic)."""
    from KGE.models.base_model.BaseModel import _Writers, WeightView
    from KGE._b200.engine import PaddedTable
    pytest.importorskip("tensorboard")
    g = torch.Generator().manual_seed(3)
    table = PaddedTable(torch.randn(1000, 50, generator=g), torch.device("cuda:0"))
    w = _Writers(str(tmp_path), with_val=False)
    seen = {}
    w.w["root"].add_histogram_raw = lambda tag, **kw2: seen.update({tag: kw2})
    w.histograms({"ent_emb": WeightView("ent_emb", table)}, 0)
    w.close()
    h = seen["ent_emb"]
    x = table.view.cpu().numpy().ravel()
    assert h["num"] == x.size and abs(h["min"] - x.min()) < 1e-6 and abs(h["max"] - x.max()) < 1e-6
    assert abs(h["sum"] - x.astype(np.float64).sum()) < 1e-3 and abs(h["sum_squares"] - (x.astype(np.float64) ** 2).sum()) < 1e-2
    want, _ = np.histogram(x, bins=30, range=(h["min"], h["max"]))
    got = np.asarray(h["bucket_counts"])
    assert got.sum() == x.size and len(h["bucket_limits"]) == 30 and np.abs(got - want).max() <= 3
    assert np.abs(np.cumsum(got) - np.cumsum(want)).max() <= 3


def test_device_batch_stream_on_the_gpu():
    """the batch pipeline lives on the device (one permutation per epoch, one gather per batch); the
    training loop reads it one batch ahead (look-ahead grouping).  Round 1 tested it on the CPU only."""
    from KGE.data_utils import DeviceBatchStream
    X = np.stack([np.arange(1000), np.arange(1000) % 7, np.arange(1000)[::-1]], 1).astype(np.int64)
    a = DeviceBatchStream(X, 128, shuffle=True, seed=3, device="cuda:0")
    b = DeviceBatchStream(X, 128, shuffle=True, seed=3, device="cuda:0")
    first = [next(a) for _ in range(20)]
    assert all(t.is_cuda and t.shape == (128, 3) and t.dtype == torch.int64 for t in first)
    epoch = torch.cat(first)[:1000]                               # the first 1000 rows are one permutation of the data
    assert torch.equal(epoch[:, 0].sort().values.cpu(), torch.arange(1000))
    assert torch.equal(epoch[:, 2], 999 - epoch[:, 0])            # rows stay intact
    b.skip(7)                                                     # resume: same draws as iterating
    assert torch.equal(next(b), first[7])
    c = DeviceBatchStream(X, 100, shuffle=False, device="cuda:0")
    assert torch.equal(next(c).cpu(), torch.from_numpy(X[:100]))


def test_epoch_permutation_and_batch_gather_match_the_spec():
    """kge_epoch_perm / kge_batch_gather (csrc/batch_stream.cu) against oracle/philox.py::epoch_permutation:
    every pass is the stable argsort of the Philox keys of (seed, pass); batches are slices of it, also across the
    boundary of two passes; int32 sources, resume by skip()."""
    from oracle import philox
    from KGE.data_utils import DeviceBatchStream
    n, B, seed = 1000, 96, 12345
    X = np.stack([np.arange(n), np.arange(n) % 11, (np.arange(n) * 7) % n], 1).astype(np.int32)
    it = DeviceBatchStream(X, B, shuffle=True, seed=seed, device="cuda:0")
    got = torch.cat([next(it) for _ in range(32)]).cpu().numpy()               # 3072 rows: three passes and a bit
    want = np.concatenate([X[philox.epoch_permutation(seed, e, n).astype(np.int64)] for e in range(4)])[:len(got)]
    assert np.array_equal(got, want.astype(np.int64))
    it2 = DeviceBatchStream(X, B, shuffle=True, seed=seed, device="cuda:0")
    it2.skip(25)                                                                # lands in the third pass
    assert np.array_equal(next(it2).cpu().numpy(), got[25 * B:26 * B])
    big = DeviceBatchStream(X[:50], 128, shuffle=True, seed=1, device="cuda:0")  # a batch longer than two passes
    rows = next(big).cpu().numpy()
    want = np.concatenate([X[:50][philox.epoch_permutation(1, e, 50).astype(np.int64)] for e in range(3)])[:128]
    assert np.array_equal(rows, want.astype(np.int64))
