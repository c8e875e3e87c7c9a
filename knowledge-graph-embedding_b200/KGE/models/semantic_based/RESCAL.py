"""RESCAL — drop-in for KGE/models/semantic_based/RESCAL.py of the reference (constructor :37-39, tables :82-103, init constraint :109-115, score :164-174, regulariser :190-198).
f = e_h^T R_r e_t with one (d x d) interaction matrix per relation; constraint=True normalises both tables
once at init and adds constraint_weight * (mean ||e||^2 + mean ||R||_F^2) (dense gradients).
Tables, init ranges and the per-batch constraint are declared here; the arithmetic is
the `RESCAL` family of the CUDA library (csrc/mat_models.cuh)."""
import math

from ...loss import SquareErrorLoss
from ...ns_strategy import UniformStrategy
from ..base_model.SemanticModel import SemanticModel
from ..._b200 import lib as _lib


class RESCAL(SemanticModel):
    MODEL = "RESCAL"

    def __init__(self, embedding_params, negative_ratio, corrupt_side, loss_fn=SquareErrorLoss(),
                 ns_strategy=UniformStrategy, constraint=True, constraint_weight=1.0, n_workers=1):
        super().__init__(embedding_params, negative_ratio, corrupt_side, loss_fn, ns_strategy, n_workers)
        self.constraint = constraint
        self.constraint_weight = constraint_weight

    def _tables(self, n_ent, n_rel):
        d, _ = self._dims()
        return {"ent_emb": ((n_ent, d), math.sqrt(6.0 / d)), "rel_inter": ((n_rel, d, d), math.sqrt(3.0 / d))}

    def _post_init(self):
        # RESCAL.py:109-115: rows to unit L2 norm, matrices to unit Frobenius norm (a matrix is one table row)
        if self.constraint:
            self._engine.constrain_table(_lib.T_ENT, _lib.CON_NORMALIZE)
            self._engine.constrain_table(_lib.T_REL_A, _lib.CON_NORMALIZE)
