"""Mapping from the reference's descriptor objects (score / loss / sampler instances,
model class names) to kge_cfg fields and table slots."""
import math
from . import lib as L

MODEL_ID = {"UM": L.UM, "SE": L.SE, "TransE": L.TRANSE, "TransH": L.TRANSH, "TransR": L.TRANSR,
            "TransD": L.TRANSD, "RotatE": L.ROTATE, "DistMult": L.DISTMULT, "RESCAL": L.RESCAL}

# model_weights key -> table slot (include/kge_abi.h)
SLOTS = {
    "UM": {"ent_emb": L.T_ENT},
    "SE": {"ent_emb": L.T_ENT, "rel_proj_h": L.T_REL_A, "rel_proj_t": L.T_REL_B},
    "TransE": {"ent_emb": L.T_ENT, "rel_emb": L.T_REL_A},
    "TransH": {"ent_emb": L.T_ENT, "rel_emb": L.T_REL_A, "rel_hyper": L.T_REL_B},
    "TransR": {"ent_emb": L.T_ENT, "rel_emb": L.T_REL_A, "rel_proj": L.T_REL_B},
    "TransD": {"ent_emb": L.T_ENT, "ent_proj": L.T_ENT_PROJ, "rel_emb": L.T_REL_A, "rel_proj": L.T_REL_B},
    "RotatE": {"ent_emb": L.T_ENT, "rel_emb": L.T_REL_A},
    "DistMult": {"ent_emb": L.T_ENT, "rel_inter": L.T_REL_A},
    "RESCAL": {"ent_emb": L.T_ENT, "rel_inter": L.T_REL_A},
}
SIDE_ID = {"h": L.SIDE_H, "t": L.SIDE_T, "h+t": L.SIDE_HT}


def score_fields(score_fn):
    """LpDistance(p) / LpDistancePow(p) / Dot (KGE/score.py:49,65,78) -> cfg fields."""
    if score_fn is None:
        return {"score": L.SCORE_DOT, "score_p": 2.0}
    name = type(score_fn).__name__
    if name == "Dot":
        return {"score": L.SCORE_DOT, "score_p": 2.0}
    if name in ("LpDistance", "LpDistancePow", "SquaredLpDistance"):
        p = float(score_fn.p)
        return {"score": L.SCORE_LP if name == "LpDistance" else L.SCORE_LP_POW,
                "score_p": p if math.isfinite(p) else float("inf")}
    raise NotImplementedError(f"score function {name} has no kernel")


def loss_fields(loss_fn):
    """KGE/loss.py classes -> cfg fields."""
    name = type(loss_fn).__name__
    if name == "PairwiseHingeLoss":
        return {"loss": L.LOSS_HINGE, "margin": float(loss_fn.margin)}
    if name == "PairwiseLogisticLoss":
        return {"loss": L.LOSS_LOGISTIC}
    if name == "BinaryCrossEntropyLoss":
        return {"loss": L.LOSS_BCE}
    if name == "SelfAdversarialNegativeSamplingLoss":
        return {"loss": L.LOSS_SELF_ADV, "margin": float(loss_fn.margin),
                "temperature": float(loss_fn.temperature)}
    if name == "SquareErrorLoss":
        return {"loss": L.LOSS_SQUARE}
    raise NotImplementedError(f"loss function {name} has no kernel")


def make_cfg(model, n_ent, n_rel, d_ent, d_rel=None, score_fn=None, loss_fn=None, negative_ratio=1,
             corrupt_side="t", constraint=True, constraint_weight=1.0, rotate_limit=None, id64=True):
    cfg = {"model": MODEL_ID[model], "n_ent": int(n_ent), "n_rel": int(n_rel), "d_ent": int(d_ent),
           "d_rel": int(d_rel if d_rel is not None else d_ent), "negative_ratio": int(negative_ratio),
           "corrupt_side": SIDE_ID[corrupt_side], "constraint": int(bool(constraint)),
           "constraint_weight": float(constraint_weight), "id_dtype": 1 if id64 else 0,
           "sampler": L.NS_UNIFORM}
    cfg.update(score_fields(score_fn))
    if loss_fn is not None:
        cfg.update(loss_fields(loss_fn))
    if model == "RotatE":
        cfg["rotate_scale"] = math.pi / float(rotate_limit)
    return cfg
