"""TranslatingModel — drop-in for KGE/models/base_model/TranslatingModel.py (:46-51):
models that score a translated head against a predicate with a pluggable `score_fn`."""
from .BaseModel import KGEModel


class TranslatingModel(KGEModel):
    HAS_SCORE_FN = True

    def __init__(self, embedding_params, negative_ratio, corrupt_side, score_fn, loss_fn, ns_strategy, n_workers):
        super().__init__(embedding_params, negative_ratio, corrupt_side, loss_fn, ns_strategy, n_workers)
        self.score_fn = score_fn
