"""Optimizer descriptor.  The reference accepts `"Adam"` or a Keras optimizer object
(BaseModel.py:243-246; tests/examples only ever pass `tf.optimizers.Adam(...)`).  The
update itself is the row kernel of the CUDA library with Keras-Adam semantics
(SURVEY.md App. B.2); this object only carries the hyper-parameters."""


class Adam:
    def __init__(self, learning_rate=0.001, beta_1=0.9, beta_2=0.999, epsilon=1e-7, amsgrad=False):
        if amsgrad:
            raise NotImplementedError("amsgrad has no kernel")
        self.learning_rate, self.beta_1, self.beta_2, self.epsilon = learning_rate, beta_1, beta_2, epsilon


def adam_fields(optimizer):
    """'Adam' | Adam | any object exposing Keras-style attributes -> kge_cfg fields."""
    if optimizer is None or optimizer == "Adam":
        optimizer = Adam()
    name = type(optimizer).__name__
    if "adam" not in name.lower():
        raise NotImplementedError(f"optimizer {name} has no kernel (Adam only)")

    def get(*names):
        for n in names:
            if hasattr(optimizer, n):
                v = getattr(optimizer, n)
                return float(v.numpy()) if hasattr(v, "numpy") else float(v)
        raise AttributeError(f"optimizer lacks {names}")
    return {"lr": get("learning_rate", "lr"), "beta1": get("beta_1", "beta1"), "beta2": get("beta_2", "beta2"),
            "eps": get("epsilon", "eps")}
