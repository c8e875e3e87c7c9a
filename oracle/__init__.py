"""CPU oracle for the KGE hot path — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package.  The product path (the ``KGE``
package under ``knowledge-graph-embedding_b200/``) never imports it and fails loudly
when the CUDA library is missing.

PARITY UNPINNED at the TensorFlow boundary: the reference is Python-on-TensorFlow 2.5
(``/root/reference/requirements.txt:7``), TensorFlow is neither installed nor
installable here, and the reference's own tests hold no golden vectors (SURVEY.md F4,
§4).  What *is* pinned:
  * the Philox4x32-10 generator against the Random123 known-answer vectors
    (``tests/test_oracle_philox.py``),
  * every invariant the reference tests assert (sign / shape / norm tolerances /
    type preservation / filtered <= raw), carried over in ``tests/test_oracle_*.py``,
  * the formulae, each cited to the reference ``file:line`` it restates.
"""
