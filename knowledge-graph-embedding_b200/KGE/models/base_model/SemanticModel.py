"""SemanticModel — drop-in for KGE/models/base_model/SemanticModel.py: models whose
interaction function is fixed (no `score_fn` argument)."""
from .BaseModel import KGEModel


class SemanticModel(KGEModel):
    HAS_SCORE_FN = False

    def __init__(self, embedding_params, negative_ratio, corrupt_side, loss_fn, ns_strategy, n_workers):
        super().__init__(embedding_params, negative_ratio, corrupt_side, loss_fn, ns_strategy, n_workers)
