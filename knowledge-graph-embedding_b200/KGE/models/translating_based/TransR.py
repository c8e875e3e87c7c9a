"""TransR — drop-in for KGE/models/translating_based/TransR.py of the reference (constructor :64-66, tables :113-130, score :178-191, constraint :207-211).
f = score_fn(clip(e_h M_r) + r, clip(e_t M_r)) with one (d_e x d_r) projection per relation,
initialised to the rectangular identity; entity and relation rows are clipped to the unit ball each batch.
Tables, init ranges and the per-batch constraint are declared here; the arithmetic is
the `TransR` family of the CUDA library (csrc/mat_models.cuh)."""
import math

from ...loss import PairwiseHingeLoss
from ...ns_strategy import UniformStrategy
from ...score import LpDistancePow
from ..base_model.TranslatingModel import TranslatingModel
from ..._b200 import lib as _lib


class TransR(TranslatingModel):
    MODEL = "TransR"

    def __init__(self, embedding_params, negative_ratio, corrupt_side, score_fn=LpDistancePow(p=2),
                 loss_fn=PairwiseHingeLoss(margin=1), ns_strategy=UniformStrategy, constraint=True, n_workers=1):
        super().__init__(embedding_params, negative_ratio, corrupt_side, score_fn, loss_fn, ns_strategy, n_workers)
        self.constraint = constraint

    def _dims(self):
        de, dr = self.embedding_params.get("ent_embedding_size"), self.embedding_params.get("rel_embedding_size")
        assert de is not None, "'ent_embedding_size' should be given in embedding_params when using TransR"
        assert dr is not None, "'rel_embedding_size' should be given in embedding_params when using TransR"
        return de, dr

    def _tables(self, n_ent, n_rel):
        de, dr = self._dims()
        return {"ent_emb": ((n_ent, de), math.sqrt(6.0 / de)), "rel_emb": ((n_rel, dr), math.sqrt(6.0 / dr)),
                "rel_proj": ((n_rel, de, dr), "eye")}
