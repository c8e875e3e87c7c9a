"""Host-side descriptor mapping, no GPU: the reference's score / loss / sampler objects and model names -> kge_cfg
fields and table slots (KGE/_b200/spec.py), and the typed sampler's pools as the CSR arrays the device sampler
reads (KGE/ns_strategy.py::_type_csr; reference: ns_strategy.py:94-151, utils.py:13-14)."""
import math

import numpy as np
import pytest

from KGE import score, loss
from KGE._b200 import lib as L, spec
from KGE.ns_strategy import _type_csr, check_typed_pools


def test_every_model_has_an_id_and_distinct_table_slots():
    assert set(spec.MODEL_ID) == {"UM", "SE", "TransE", "TransH", "TransR", "TransD", "RotatE", "DistMult", "RESCAL"}
    assert len(set(spec.MODEL_ID.values())) == 9
    for model, slots in spec.SLOTS.items():
        assert slots["ent_emb"] == L.T_ENT and len(set(slots.values())) == len(slots), model
        assert all(0 <= s < 4 for s in slots.values())          # KGE_MAX_TABLES


@pytest.mark.parametrize("fn,want", [
    (None, (L.SCORE_DOT, 2.0)), (score.Dot(), (L.SCORE_DOT, 2.0)),
    (score.LpDistance(p=1), (L.SCORE_LP, 1.0)), (score.LpDistance(p=2), (L.SCORE_LP, 2.0)),
    (score.LpDistance(p=3), (L.SCORE_LP, 3.0)), (score.LpDistance(p=math.inf), (L.SCORE_LP, math.inf)),
    (score.LpDistancePow(p=2), (L.SCORE_LP_POW, 2.0))])
def test_score_objects_map_to_cfg_fields(fn, want):
    f = spec.score_fields(fn)
    assert (f["score"], f["score_p"]) == want


def test_loss_objects_map_to_cfg_fields():
    assert spec.loss_fields(loss.PairwiseHingeLoss(margin=2.5)) == {"loss": L.LOSS_HINGE, "margin": 2.5}
    assert spec.loss_fields(loss.PairwiseLogisticLoss()) == {"loss": L.LOSS_LOGISTIC}
    assert spec.loss_fields(loss.BinaryCrossEntropyLoss()) == {"loss": L.LOSS_BCE}
    assert spec.loss_fields(loss.SquareErrorLoss()) == {"loss": L.LOSS_SQUARE}
    sa = spec.loss_fields(loss.SelfAdversarialNegativeSamplingLoss(margin=3.0, temperature=0.5))
    assert sa == {"loss": L.LOSS_SELF_ADV, "margin": 3.0, "temperature": 0.5}

    class Unknown:
        pass
    with pytest.raises(NotImplementedError):
        spec.loss_fields(Unknown())
    with pytest.raises(NotImplementedError):
        spec.score_fields(Unknown())


def test_make_cfg_fills_a_struct_the_library_accepts():
    import ctypes as C
    kw = spec.make_cfg("TransD", 100, 7, 20, 12, score_fn=score.LpDistancePow(p=2), loss_fn=loss.PairwiseHingeLoss(margin=1.0),
                       negative_ratio=5, corrupt_side="h+t", constraint=True, constraint_weight=0.5)
    assert (kw["d_ent"], kw["d_rel"], kw["corrupt_side"], kw["negative_ratio"]) == (20, 12, L.SIDE_HT, 5)
    assert kw["id_dtype"] == 1 and kw["sampler"] == L.NS_UNIFORM and kw["constraint"] == 1
    cfg = L.Cfg(abi_version=L.KGE_ABI_VERSION, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-7, clip_norm=5.0, **kw)
    n = C.c_size_t()
    # check_cfg passes (the ranking workspace needs no device; a training workspace asks CUB, i.e. a GPU)
    assert L.load().kge_workspace_bytes(C.byref(cfg), 0, 64, C.byref(n)) == 0 and n.value > 0
    bad = L.Cfg(abi_version=L.KGE_ABI_VERSION, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-7, clip_norm=5.0, **dict(kw, model=99))
    assert L.load().kge_workspace_bytes(C.byref(bad), 0, 64, C.byref(n)) == -1
    assert spec.make_cfg("TransE", 10, 2, 8)["d_rel"] == 8                       # d_rel defaults to d_ent
    rot = spec.make_cfg("RotatE", 10, 2, 8, rotate_limit=0.25)
    assert rot["rotate_scale"] == math.pi / 0.25
    with pytest.raises(KeyError):
        spec.make_cfg("ComplEx", 10, 2, 8)
    with pytest.raises(KeyError):
        spec.make_cfg("TransE", 10, 2, 8, corrupt_side="both")


def test_typed_pools_as_csr():
    ind2type = ["person", "city", "person", "country", "city", "person"]
    ent_type, off, members, pos = _type_csr(ind2type)
    names = sorted(set(ind2type))
    for e, t in enumerate(ind2type):
        ti = ent_type[e]
        pool = members[off[ti]:off[ti + 1]]
        assert names[ti] == t and e in pool
        assert sorted(pool.tolist()) == [i for i, tt in enumerate(ind2type) if tt == t]
        assert pool[pos[e]] == e                                  # where the entity sits in its own pool (self-skip)
    assert off[0] == 0 and off[-1] == len(ind2type)
    with pytest.raises(ValueError):                                # 'country' has one member: nothing else to draw
        check_typed_pools(ent_type, off, pos)
    ok_type, ok_off, _, ok_pos = _type_csr(["a", "b", "a", "b"])
    check_typed_pools(ok_type, ok_off, ok_pos)
    # pools given by the caller (metadata['type2inds']) may list entities of OTHER types: those have no own position
    t2i = {"a": np.array([0, 2, 3]), "b": np.array([1, 3])}
    ent_type, off, members, pos = _type_csr(["a", "b", "a", "b"], t2i)
    assert members.tolist() == [0, 2, 3, 1, 3] and off.tolist() == [0, 3, 5]
    assert pos.tolist() == [0, 0, 1, 1]                            # entity 3 is of type b: its position in b's pool


def test_optimizer_descriptors_map_to_cfg_fields():
    """'Adam' / 'SGD' / 'Adagrad', the KGE.optimizers objects and any Keras-style object (attributes that may be
    tf.Variables: `.numpy()`) -> kge_cfg fields; what has no kernel raises (BaseModel.py:243-246 accepts any optimizer)."""
    from KGE import optimizers as O

    assert O.adam_fields(None) == O.adam_fields("Adam") == {"optimizer": L.OPTIM_ADAM, "lr": 0.001, "beta1": 0.9, "beta2": 0.999, "eps": 1e-7}
    assert O.adam_fields(O.Adam(learning_rate=0.01, beta_1=0.8, beta_2=0.99, epsilon=1e-6)) == {
        "optimizer": L.OPTIM_ADAM, "lr": 0.01, "beta1": 0.8, "beta2": 0.99, "eps": 1e-6}
    assert O.adam_fields("sgd") == {"optimizer": L.OPTIM_SGD, "lr": 0.01}
    assert O.adam_fields(O.Adagrad(learning_rate=0.05, initial_accumulator_value=0.2)) == {
        "optimizer": L.OPTIM_ADAGRAD, "lr": 0.05, "eps": 1e-7, "acc_init": 0.2}

    class Var:                       # a tf.Variable-like hyper-parameter
        def __init__(self, v):
            self.v = v

        def numpy(self):
            return self.v

    class KerasAdam:                 # name contains "adam": picked up by its attributes
        learning_rate, beta_1, beta_2, epsilon = Var(0.002), 0.85, 0.995, 1e-8
    got = O.adam_fields(KerasAdam())
    assert got == {"optimizer": L.OPTIM_ADAM, "lr": 0.002, "beta1": 0.85, "beta2": 0.995, "eps": 1e-8}

    class SGD:
        learning_rate, momentum = 0.1, 0.9
    for bad in (lambda: O.adam_fields("RMSprop"), lambda: O.adam_fields(SGD()), lambda: O.SGD(momentum=0.5),
                lambda: O.Adam(amsgrad=True), lambda: O.adam_fields(type("Nadam", (), {})())):
        with pytest.raises(NotImplementedError):
            bad()
