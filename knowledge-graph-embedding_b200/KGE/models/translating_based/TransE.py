"""TransE — drop-in for KGE/models/translating_based/TransE.py of the reference (constructor :48-50, tables :94-103, score :151-155, constraint :171-174).
f(h, r, t) = score_fn(e_h + r, e_t); entities L2-normalised every batch, relations once at init.
Tables, init ranges and the per-batch constraint are declared here; the arithmetic is
the `TransE` family of the CUDA library (csrc/models.cuh)."""
import math

from ...loss import PairwiseHingeLoss
from ...ns_strategy import UniformStrategy
from ...score import LpDistance
from ..base_model.TranslatingModel import TranslatingModel
from ..._b200 import lib as _lib


class TransE(TranslatingModel):
    MODEL = "TransE"

    def __init__(self, embedding_params, negative_ratio, corrupt_side, score_fn=LpDistance(p=2),
                 loss_fn=PairwiseHingeLoss(margin=1), ns_strategy=UniformStrategy, constraint=True, n_workers=1):
        super().__init__(embedding_params, negative_ratio, corrupt_side, score_fn, loss_fn, ns_strategy, n_workers)
        self.constraint = constraint

    def _tables(self, n_ent, n_rel):
        d, _ = self._dims()
        lim = 6.0 / math.sqrt(d)
        return {"ent_emb": ((n_ent, d), lim), "rel_emb": ((n_rel, d), lim)}

    def _post_init(self):
        # TransE.py:108-109 (works with model_weights_initial too: SURVEY.md C.2 fixed)
        if self.constraint:
            for eng in self._train_engines():        # the relation table is replicated on every rank
                eng.constrain_table(_lib.T_REL_A, _lib.CON_NORMALIZE)
