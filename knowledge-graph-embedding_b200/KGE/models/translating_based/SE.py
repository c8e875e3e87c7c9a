"""SE — drop-in for KGE/models/translating_based/SE.py of the reference (constructor :45-47, tables :90-107, score :155-161, constraint :177-180).
f = score_fn(R^h_r e_h, R^t_r e_t) with two (d x d) matrices per relation; entities L2-normalised each batch.
Tables, init ranges and the per-batch constraint are declared here; the arithmetic is
the `SE` family of the CUDA library (csrc/mat_models.cuh)."""
import math

from ...loss import PairwiseHingeLoss
from ...ns_strategy import UniformStrategy
from ...score import LpDistance
from ..base_model.TranslatingModel import TranslatingModel
from ..._b200 import lib as _lib


class SE(TranslatingModel):
    MODEL = "SE"

    def __init__(self, embedding_params, negative_ratio, corrupt_side, score_fn=LpDistance(p=1),
                 loss_fn=PairwiseHingeLoss(margin=1), ns_strategy=UniformStrategy, constraint=True, n_workers=1):
        super().__init__(embedding_params, negative_ratio, corrupt_side, score_fn, loss_fn, ns_strategy, n_workers)
        self.constraint = constraint

    def _tables(self, n_ent, n_rel):
        d, _ = self._dims()
        lm = math.sqrt(3.0 / d)
        return {"ent_emb": ((n_ent, d), math.sqrt(6.0 / d)), "rel_proj_h": ((n_rel, d, d), lm),
                "rel_proj_t": ((n_rel, d, d), lm)}
