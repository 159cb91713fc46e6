"""CPU oracle for the ternary-Linear hot path.  TEST INFRASTRUCTURE ONLY.

This package restates, on the CPU, the algorithms of the reference
(BenCrulis/ternary-vllm) for the one path this repo accelerates.  It is imported
only by ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs.  Nothing under
``ternary_vllm_b200/`` may import it: the product path is CUDA-only and fails
loudly when the extension is missing.

Pinning status
--------------
* ``ref_modules`` (PyTorch fake-quant layers, ``binary/modules.py`` and
  ``binary/utils.py``): PINNED.  Checked against the reference imported live in
  the build container (``tests/test_oracle_vs_reference.py``) and against the
  committed fixtures ``tests/golden/*.npz`` that ``oracle/make_golden.py``
  generated from the live reference.
* ``ref_packing`` formats A/B/C and the packed layers' forward
  (``impl/tf/moondream/ternary.py``): PINNED.  TensorFlow is absent, so
  ``oracle/make_golden_packing.py`` runs the reference's own source text of
  ``pack_ternary``, ``pack_ternary2``, the unpackers, ``ScaledTernaryPackedMM.pack_fn``
  / ``.call`` and ``ScaledTernaryTf.call`` (and the reference's own
  ``TernaryCase`` test) with ``oracle/tf_stub.py`` bound to ``tf``; the outputs are
  committed as ``tests/golden/packing_vectors.npz`` and re-derived live in the build
  container by ``tests/test_oracle_packing.py``.
* ``larq_compute_engine.ops.{ternary_matmul,pack_fn,unpack_ternary}`` (author's
  private LCE fork, no version pin, SOURCES ABSENT): nothing to execute.  The contract
  is tied to in-tree code instead: the ``[N, ceil(K/4)]`` weight shape the ops consume
  (``ternary.py:264-276,316-328``) is the shape ``ScaledTernaryPackedMM.pack_fn``
  produces (format C, pinned above), and the fused op's arithmetic is spelled out by
  ``ScaledTernaryTf.call`` / the commented fallback at ``ternary.py:289-297`` (pinned
  above).
"""
