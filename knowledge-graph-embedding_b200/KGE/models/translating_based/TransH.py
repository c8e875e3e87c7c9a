"""TransH — drop-in for KGE/models/translating_based/TransH.py of the reference (constructor :64-66, tables :112-125, score :173-185, constraint :202-211).
f = score_fn(e_h - (w.e_h)w + r, e_t - (w.e_t)w); constraint=True normalises the hyperplanes each
batch and adds the soft scale / orthogonality penalties (dense gradients) weighted by constraint_weight.
Tables, init ranges and the per-batch constraint are declared here; the arithmetic is
the `TransH` family of the CUDA library (csrc/models.cuh)."""
import math

from ...loss import PairwiseHingeLoss
from ...ns_strategy import UniformStrategy
from ...score import LpDistancePow
from ..base_model.TranslatingModel import TranslatingModel
from ..._b200 import lib as _lib


class TransH(TranslatingModel):
    MODEL = "TransH"

    def __init__(self, embedding_params, negative_ratio, corrupt_side, score_fn=LpDistancePow(p=2),
                 loss_fn=PairwiseHingeLoss(margin=1), ns_strategy=UniformStrategy, constraint=True,
                 constraint_weight=1.0, n_workers=1):
        super().__init__(embedding_params, negative_ratio, corrupt_side, score_fn, loss_fn, ns_strategy, n_workers)
        self.constraint = constraint
        self.constraint_weight = constraint_weight

    def _tables(self, n_ent, n_rel):
        d, _ = self._dims()
        lim = math.sqrt(6.0 / d)
        return {"ent_emb": ((n_ent, d), lim), "rel_emb": ((n_rel, d), lim), "rel_hyper": ((n_rel, d), lim)}
