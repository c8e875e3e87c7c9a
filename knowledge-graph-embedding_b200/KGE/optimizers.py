"""Optimizer descriptors.  The reference accepts `"Adam"` or any Keras optimizer object
(BaseModel.py:243-246; its tests / examples pass `tf.optimizers.Adam(...)`).  The update itself is the
row kernel of the CUDA library with the TF 2.5 OptimizerV2 sparse-apply semantics (SURVEY.md App. B.2,
include/kge_abi.h kge_opt_kind); these objects only carry the hyper-parameters."""
from ._b200 import lib as _L


class Adam:
    def __init__(self, learning_rate=0.001, beta_1=0.9, beta_2=0.999, epsilon=1e-7, amsgrad=False):
        if amsgrad:
            raise NotImplementedError("amsgrad has no kernel")
        self.learning_rate, self.beta_1, self.beta_2, self.epsilon = learning_rate, beta_1, beta_2, epsilon


class SGD:
    """tf.optimizers.SGD: w -= lr * g on the looked-up rows (momentum has no kernel)."""

    def __init__(self, learning_rate=0.01, momentum=0.0, nesterov=False):
        if momentum or nesterov:
            raise NotImplementedError("SGD momentum / nesterov has no kernel")
        self.learning_rate, self.momentum = learning_rate, 0.0


class Adagrad:
    """tf.optimizers.Adagrad: acc += g^2; w -= lr * g / (sqrt(acc) + eps) on the looked-up rows."""

    def __init__(self, learning_rate=0.001, initial_accumulator_value=0.1, epsilon=1e-7):
        self.learning_rate, self.initial_accumulator_value, self.epsilon = learning_rate, initial_accumulator_value, epsilon


def _get(optimizer, *names, default=None):
    for n in names:
        if hasattr(optimizer, n):
            v = getattr(optimizer, n)
            return float(v.numpy()) if hasattr(v, "numpy") else float(v)
    if default is not None:
        return float(default)
    raise AttributeError(f"optimizer lacks {names}")


def adam_fields(optimizer):
    """'Adam' | 'SGD' | 'Adagrad' | any object exposing Keras-style attributes -> kge_cfg fields
    (+ 'acc_init' for the engine: initial value of the Adagrad accumulator slot)."""
    if optimizer is None:
        optimizer = "Adam"
    if isinstance(optimizer, str):
        known = {"adam": Adam, "sgd": SGD, "adagrad": Adagrad}
        if optimizer.lower() not in known:
            raise NotImplementedError(f"optimizer {optimizer} has no kernel (Adam, SGD, Adagrad)")
        optimizer = known[optimizer.lower()]()
    name = type(optimizer).__name__.lower()
    if "adam" in name and "adamax" not in name and "nadam" not in name:
        return {"optimizer": _L.OPTIM_ADAM, "lr": _get(optimizer, "learning_rate", "lr"),
                "beta1": _get(optimizer, "beta_1", "beta1"), "beta2": _get(optimizer, "beta_2", "beta2"),
                "eps": _get(optimizer, "epsilon", "eps")}
    if "sgd" in name:
        if _get(optimizer, "momentum", default=0.0) != 0.0 or getattr(optimizer, "nesterov", False):
            raise NotImplementedError("SGD momentum / nesterov has no kernel")
        return {"optimizer": _L.OPTIM_SGD, "lr": _get(optimizer, "learning_rate", "lr")}
    if "adagrad" in name:
        return {"optimizer": _L.OPTIM_ADAGRAD, "lr": _get(optimizer, "learning_rate", "lr"),
                "eps": _get(optimizer, "epsilon", "eps", default=1e-7),
                "acc_init": _get(optimizer, "initial_accumulator_value", default=0.1)}
    raise NotImplementedError(f"optimizer {type(optimizer).__name__} has no kernel (Adam, SGD, Adagrad)")
