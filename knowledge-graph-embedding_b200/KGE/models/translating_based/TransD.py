"""TransD — drop-in for KGE/models/translating_based/TransD.py of the reference (constructor :74-76, tables :123-143, score :194-222, constraint :238-242).
f = score_fn(clip((r_p e_p^T + I) e_h) + r, clip((r_p e_p^T + I) e_t)); evaluated in the rank-1 form
(e_p . e) r_p + e[:d_r], never building the (d_r x d_e) matrices.
Tables, init ranges and the per-batch constraint are declared here; the arithmetic is
the `TransD` family of the CUDA library (csrc/models.cuh)."""
import math

from ...loss import PairwiseHingeLoss
from ...ns_strategy import UniformStrategy
from ...score import LpDistancePow
from ..base_model.TranslatingModel import TranslatingModel
from ..._b200 import lib as _lib


class TransD(TranslatingModel):
    MODEL = "TransD"

    def __init__(self, embedding_params, negative_ratio, corrupt_side, score_fn=LpDistancePow(p=2),
                 loss_fn=PairwiseHingeLoss(margin=1), ns_strategy=UniformStrategy, constraint=True, n_workers=1):
        super().__init__(embedding_params, negative_ratio, corrupt_side, score_fn, loss_fn, ns_strategy, n_workers)
        self.constraint = constraint

    def _dims(self):
        de, dr = self.embedding_params.get("ent_embedding_size"), self.embedding_params.get("rel_embedding_size")
        assert de is not None, "'ent_embedding_size' should be given in embedding_params when using TransD"
        assert dr is not None, "'rel_embedding_size' should be given in embedding_params when using TransD"
        return de, dr

    def _tables(self, n_ent, n_rel):
        de, dr = self._dims()
        le, lr = math.sqrt(6.0 / de), math.sqrt(6.0 / dr)
        return {"ent_emb": ((n_ent, de), le), "ent_proj": ((n_ent, de), le),
                "rel_emb": ((n_rel, dr), lr), "rel_proj": ((n_rel, dr), lr)}
