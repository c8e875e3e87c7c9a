"""KGE — drop-in Python API of melissakou/knowledge-graph-embedding whose batch hot
path (training step, scoring, filtered ranking) runs as sm_100a CUDA kernels through
libkge_b200.so (include/kge_abi.h).  No TensorFlow, no CPU fallback."""
from . import constraint, loss, metrics, ns_strategy, optimizers, score  # noqa: F401
