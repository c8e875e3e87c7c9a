"""CPU oracle for the KGE hot path — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package.  The product path (the ``KGE``
package under ``knowledge-graph-embedding_b200/``) never imports it and fails loudly
when the CUDA library is missing.

PARITY: pinned to the reference's own SOURCE, unpinned at the TensorFlow-kernel boundary.
The reference is Python-on-TensorFlow 2.5 (``/root/reference/requirements.txt:7``);
TensorFlow is neither installed nor installable here and the reference's own tests hold
no golden vectors.  So the reference's unmodified Python
(``/root/reference/KGE``: scores, losses, constraints, all nine models, ``train()``,
``__run_single_batch``, ``get_rank``, ``evaluate``) is executed over
``oracle/tf_shim/tensorflow`` — a torch-backed stand-in for the ~50 ``tf.*`` entry points
it touches, with the TF semantics of SURVEY.md Appendix B — by
``tests/golden/make_golden.py``; the outputs are committed under ``tests/golden/`` and
``tests/test_golden_oracle.py`` checks this oracle against them (17 model cases + a full
2-epoch train loop).  What is NOT pinned: TensorFlow's own kernels/rounding and its RNG
streams (the shim is not TensorFlow).  Also pinned:
  * the Philox4x32-10 generator against the Random123 known-answer vectors
    (``tests/test_oracle_philox.py``),
  * every invariant the reference tests assert (sign / shape / norm tolerances /
    type preservation / filtered <= raw), carried over in ``tests/test_oracle_*.py``,
  * the formulae, each cited to the reference ``file:line`` it restates.
"""
