"""UM — drop-in for KGE/models/translating_based/UM.py of the reference (constructor :42-44, tables :87-92, score :134-137, constraint :153-156).
f(h, t) = score_fn(e_h, e_t); entity rows are L2-normalised at the start of every batch.
Tables, init ranges and the per-batch constraint are declared here; the arithmetic is
the `UM` family of the CUDA library (csrc/models.cuh)."""
import math

from ...loss import PairwiseHingeLoss
from ...ns_strategy import UniformStrategy
from ...score import LpDistancePow
from ..base_model.TranslatingModel import TranslatingModel
from ..._b200 import lib as _lib


class UM(TranslatingModel):
    MODEL = "UM"

    def __init__(self, embedding_params, negative_ratio, corrupt_side, score_fn=LpDistancePow(p=2),
                 loss_fn=PairwiseHingeLoss(margin=1), ns_strategy=UniformStrategy, constraint=True, n_workers=1):
        super().__init__(embedding_params, negative_ratio, corrupt_side, score_fn, loss_fn, ns_strategy, n_workers)
        self.constraint = constraint

    def _tables(self, n_ent, n_rel):
        d, _ = self._dims()
        return {"ent_emb": ((n_ent, d), math.sqrt(6.0 / d))}
