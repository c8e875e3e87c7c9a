"""RotatE — drop-in for KGE/models/translating_based/RotatE.py of the reference (constructor :46-49, tables :88-105, score :150-165).
f = score_fn(h o exp(i theta_r), t) on complex entity rows (re, im interleaved), theta = rel_emb / limit * pi,
limit = (margin + 2) / d; no constraint.
Tables, init ranges and the per-batch constraint are declared here; the arithmetic is
the `RotatE` family of the CUDA library (csrc/models.cuh)."""
import math

from ...loss import SelfAdversarialNegativeSamplingLoss
from ...ns_strategy import UniformStrategy
from ...score import LpDistance
from ..base_model.TranslatingModel import TranslatingModel
from ..._b200 import lib as _lib


class RotatE(TranslatingModel):
    MODEL = "RotatE"

    def __init__(self, embedding_params, negative_ratio, corrupt_side, score_fn=LpDistance(p=1),
                 loss_fn=SelfAdversarialNegativeSamplingLoss(margin=3, temperature=1), ns_strategy=UniformStrategy,
                 n_workers=1):
        super().__init__(embedding_params, negative_ratio, corrupt_side, score_fn, loss_fn, ns_strategy, n_workers)

    def _rotate_limit(self):
        margin = self.loss_fn.margin if hasattr(self.loss_fn, "margin") else 6.0
        self.limit = (margin + 2.0) / self._dims()[0]
        return self.limit

    def _tables(self, n_ent, n_rel):
        d, _ = self._dims()
        lim = self._rotate_limit()
        return {"ent_emb": ((n_ent, d, 2), lim), "rel_emb": ((n_rel, d), lim)}
