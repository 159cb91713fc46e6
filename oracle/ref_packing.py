"""Numpy restatement of the reference's 2-bit ternary pack formats.  TEST ORACLE ONLY.

Follows ``/root/reference/impl/tf/moondream/ternary.py``:

* format A  ``pack_ternary`` (:57-78) / ``unpack_ternary_iterative`` (:101-110)
* format B  ``pack_ternary2`` (:114-133) / ``unpack_ternary_iterative2`` (:136-156)
* format C  ``ScaledTernaryPackedMM.pack_fn`` (:374-398) — same ``[N, ceil(K/4)]``
  shape as the ``pw`` the LCE-fork ops consume (:264-276, :316-328).

Parity: PINNED.  ``oracle/make_golden_packing.py`` executes the reference's own source of
these functions (taken from ``ternary.py`` with ``ast``; ``oracle/tf_stub.py`` stands in for
the absent TensorFlow) and commits its outputs as ``tests/golden/packing_vectors.npz``;
``tests/test_oracle_packing.py`` holds this restatement to those bytes (ragged lengths, the
full alphabet, all 256 byte values through the unpackers), to the round trip the reference
tests (``ternary_tests.py:17-38``, also executed from source) and to hand-derived
known-answer bytes.  One reference quirk: for K % 4 != 0 the 2-D ``pack_fn`` raises
(``tf.concat`` rank mismatch, ``ternary.py:380``); padding such rows with code 1 is an
extension defined here.
"""
from __future__ import annotations

import numpy as np


def _pad4_flat(t: np.ndarray) -> np.ndarray:
    """ternary.py:58-63 / :116-121 — pad a 1-D vector with zeros to a multiple of 4."""
    t = np.asarray(t)
    to_add = (4 - t.shape[-1] % 4) % 4
    if to_add:
        t = np.concatenate([t, np.zeros((to_add,), dtype=t.dtype)], axis=-1)
    return t


def pack_A(t: np.ndarray) -> np.ndarray:
    """Format A, ternary.py:57-78.  1-D input.  Planar: byte j, bits [2i,2i+1] hold
    element ``i*S + j`` (S = padded_len/4); bit0 = (t==+1), bit1 = (t==-1)."""
    t = _pad4_flat(np.asarray(t).reshape(-1))
    S = t.shape[0] // 4
    packed = np.zeros((S,), dtype=np.uint8)
    for i in range(4):
        s = t[i * S:(i + 1) * S]
        val = (s == 1).astype(np.uint8) ^ ((s == -1).astype(np.uint8) << 1)
        packed |= (val << (2 * i)).astype(np.uint8)
    return packed


def unpack_A(packed: np.ndarray, dtype=np.float32) -> np.ndarray:
    """ternary.py:101-110.  Returns the padded-length vector (caller slices).
    Code 0b11 decodes to 1 - 1 = 0."""
    packed = np.asarray(packed, dtype=np.uint8)
    res = []
    for i in range(4):
        shifted = packed >> np.uint8(2 * i)
        pos = shifted & np.uint8(1)
        neg = (shifted & np.uint8(3)) >> np.uint8(1)
        res.append(pos.astype(dtype) - neg.astype(dtype))
    return np.concatenate(res, axis=-1)


def pack_B(t: np.ndarray) -> np.ndarray:
    """Format B, ternary.py:114-133.  Flattens, pads with 0, planar like A, code = t+1."""
    t = _pad4_flat(np.asarray(t).reshape(-1))
    S = t.shape[0] // 4
    packed = np.zeros((S,), dtype=np.uint8)
    for i in range(4):
        s = t[i * S:(i + 1) * S]
        packed |= ((s + 1).astype(np.uint8) << np.uint8(2 * i)).astype(np.uint8)
    return packed


def unpack_B(packed: np.ndarray, dtype=np.float32) -> np.ndarray:
    """ternary.py:136-156.  Code 0b11 decodes to +2 (the reference does not guard it)."""
    packed = np.asarray(packed, dtype=np.uint8)
    outs = []
    for i in range(4):
        f = (packed >> np.uint8(2 * i)) & np.uint8(3)
        outs.append(f.astype(dtype) - dtype(1))
    return np.concatenate(outs, axis=-1)


def pack_C(w: np.ndarray) -> np.ndarray:
    """Format C, ternary.py:374-398.  ``w`` is [N, K] in {-1,0,1}.  Rows are padded to a
    multiple of 4 with 0 (code 1); byte j of row n, bits [2i,2i+1] = w[n, 4j+i] + 1."""
    w = np.asarray(w)
    assert w.ndim == 2
    N, K = w.shape
    to_add = (4 - K % 4) % 4
    if to_add:
        w = np.concatenate([w, np.zeros((N, to_add), dtype=w.dtype)], axis=-1)
    codes = (w + 1).astype(np.uint8).reshape(N, -1, 4)
    packed = np.zeros(codes.shape[:2], dtype=np.uint8)
    for i in range(4):
        packed |= (codes[:, :, i] << np.uint8(2 * i)).astype(np.uint8)
    return packed


def unpack_C(packed: np.ndarray, K: int, dtype=np.int8) -> np.ndarray:
    """Inverse of :func:`pack_C` (``ops.unpack_ternary(pw, orig_dim)``, ternary.py:334;
    decode rule spelled out at ternary.py:505-519): field - 1, first K columns."""
    packed = np.asarray(packed, dtype=np.uint8)
    N = packed.shape[0]
    out = np.empty((N, packed.shape[1], 4), dtype=np.int16)
    for i in range(4):
        out[:, :, i] = ((packed >> np.uint8(2 * i)) & np.uint8(3)).astype(np.int16) - 1
    return out.reshape(N, -1)[:, :K].astype(dtype)


def packed_mm_C(x: np.ndarray, packed: np.ndarray) -> np.ndarray:
    """The strided-activation x bit-plane formulation, ternary.py:505-533:
    ``out = sum_i x[..., i::4] @ plane_i.T`` (x zero-padded to a multiple of 4)."""
    x = np.asarray(x)
    K = x.shape[-1]
    to_pad = (4 - K % 4) % 4
    if to_pad:
        x = np.concatenate([x, np.zeros(x.shape[:-1] + (to_pad,), dtype=x.dtype)], axis=-1)
    out = 0
    for i in range(4):
        plane = ((packed >> np.uint8(2 * i)) & np.uint8(3)).astype(x.dtype) - x.dtype.type(1)
        out = out + x[..., i::4] @ plane.T
    return out


# Known-answer vectors: t = [1,-1,0,1, 0,0,-1,1]
KAT_T = np.array([1, -1, 0, 1, 0, 0, -1, 1], dtype=np.int32)
KAT_A = bytes([0x81, 0x46])
KAT_B = bytes([0x16, 0x98])
KAT_C = bytes([0x92, 0x85])
