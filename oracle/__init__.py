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
* ``ref_packing`` formats A/B/C (``impl/tf/moondream/ternary.py``): pinned
  against the reference's only test for them (round trip on a length-5031
  vector, ``ternary_tests.py:17-38``) plus hand-derived known-answer bytes.  The
  TensorFlow code itself cannot be run here (TF absent), so byte-level parity
  with the TF functions is by restatement: "parity unpinned" beyond the round
  trip and the known-answer vectors.
* ``larq_compute_engine.ops.{ternary_matmul,pack_fn,unpack_ternary}`` (author's
  private LCE fork, no version pin, sources absent): PARITY UNPINNED.  The
  contract is restated from the call sites (``ternary.py:279-297,334``,
  ``conversion.py:69-72``).
"""
