"""DistMult — drop-in for KGE/models/semantic_based/DistMult.py of the reference (constructor :38-40, tables :83-97, score :142-146, constraint :162-165).
f = sum(e_h * r * e_t); constraint=True normalises entities each batch and adds
constraint_weight * mean_batch ||r||^2.
Tables, init ranges and the per-batch constraint are declared here; the arithmetic is
the `DistMult` family of the CUDA library (csrc/models.cuh)."""
import math

from ...loss import PairwiseHingeLoss
from ...ns_strategy import UniformStrategy
from ..base_model.SemanticModel import SemanticModel
from ..._b200 import lib as _lib


class DistMult(SemanticModel):
    MODEL = "DistMult"

    def __init__(self, embedding_params, negative_ratio, corrupt_side, loss_fn=PairwiseHingeLoss(margin=1),
                 ns_strategy=UniformStrategy, constraint=True, constraint_weight=1.0, n_workers=1):
        super().__init__(embedding_params, negative_ratio, corrupt_side, loss_fn, ns_strategy, n_workers)
        self.constraint = constraint
        self.constraint_weight = constraint_weight

    def _tables(self, n_ent, n_rel):
        d, _ = self._dims()
        lim = math.sqrt(6.0 / d)
        return {"ent_emb": ((n_ent, d), lim), "rel_inter": ((n_rel, d), lim)}
