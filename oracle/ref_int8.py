"""Numpy restatement of the integer path the B200 kernels implement.  TEST ORACLE ONLY.

The reference has NO activation quantisation (binary/modules.py:101-109 multiplies float
activations with the fake-quantised weights).  BASELINE.json's north_star asks for per-token
int8 activation quantisation fused into the kernels, so this file defines that arithmetic
exactly (it is the specification the CUDA kernels are checked against, bit for bit on the
integer accumulators) and ties it back to the reference: feeding the de-quantised
activations ``q_x * s_x`` to the reference layer must reproduce the fused output within 1e-3
(tests/test_gpu_*.py, SURVEY.md §7 "Parity definition").

  s_x[m]  = max_k |x[m,k]| / 127            fp32 division; 1.0 for an all-zero row
  inv[m]  = 127 / max_k |x[m,k]|            fp32 division; 1.0 for an all-zero row
  q[m,k]  = clip(rint(x[m,k] * inv[m]), -127, 127)        fp32 multiply, round-half-even
  acc     = q (int64) @ codes(W)^T (int64)                 codes per binary/modules.py:21-26
  y       = clip((float32(acc) * s_x) * scale + bias, +-clip_val)   all fp32
"""
from __future__ import annotations

import numpy as np


def quant_act(x: np.ndarray):
    """x [M,K] float32 (bf16/fp16 inputs are widened exactly first) -> (q int8, s_x f32, sum i32)."""
    x = np.asarray(x, dtype=np.float32)
    amax = np.abs(x).max(axis=1) if x.shape[1] else np.zeros(x.shape[0], np.float32)
    safe = np.where(amax > 0, amax, np.float32(1.0)).astype(np.float32)
    s = np.where(amax > 0, (safe / np.float32(127.0)).astype(np.float32), np.float32(1.0)).astype(np.float32)
    inv = np.where(amax > 0, (np.float32(127.0) / safe).astype(np.float32), np.float32(1.0)).astype(np.float32)
    q = np.rint((x * inv[:, None]).astype(np.float32))
    q = np.clip(q, -127, 127).astype(np.int8)
    return q, s, q.astype(np.int32).sum(axis=1).astype(np.int32)


def kperm_index(K_padded: int) -> np.ndarray:
    """Position p of the tcgen05 operand order holds element kperm_index[p]: within every group of
    16, position 4i+b holds element 4b+i (ternary_b200.h, tb200_quant_act_i8 kperm=1)."""
    assert K_padded % 16 == 0
    p = np.arange(K_padded)
    g, r = p // 16, p % 16
    return g * 16 + 4 * (r % 4) + r // 4


def split_layout_index(K_padded: int) -> np.ndarray:
    """tb200_quant_act_i8 layout 3 (TB200_QLAYOUT_GEMM_SPLIT, operand order of the field-split 32-token tcgen05 kernel):
    inside every 256-element block, element 16 j + 4 b + f (f = 2-bit field of packed byte b of packed word j) sits at
    position 128 (f & 1) + 8 j + 4 (f >> 1) + b.  Returns idx with: position p holds element idx[p]."""
    assert K_padded % 256 == 0
    k = np.arange(256)
    j, b, f = k // 16, (k % 16) // 4, k % 4
    pos = 128 * (f & 1) + 8 * j + 4 * (f >> 1) + b
    inv = np.empty(256, dtype=np.int64)
    inv[pos] = k
    return (np.arange(K_padded // 256)[:, None] * 256 + inv[None, :]).reshape(-1)


def gemv_layout_index(K: int) -> np.ndarray:
    """tb200_quant_act_i8 layout 2 (the decode GEMV's shared-memory image): byte position p of the
    row holds element gemv_layout_index[p] of the zero-padded activation row (ldq = 4096*ceil(K/4096)).
    16-element group g (kperm order inside) lands at uint4 slot ((c>>5)*4 + j)*32 + (c&31), c=g>>2, j=g&3."""
    ldq = (K + 4095) // 4096 * 4096
    src = kperm_index(ldq).reshape(-1, 16)          # group g -> its 16 source elements in kperm order
    out = np.empty((ldq // 16, 16), dtype=np.int64)
    g = np.arange(ldq // 16)
    c, j = g >> 2, g & 3
    out[((c >> 5) * 4 + j) * 32 + (c & 31)] = src
    return out.reshape(-1)


def int_acc(q_x: np.ndarray, codes: np.ndarray) -> np.ndarray:
    """Exact integer accumulators: q_x [M,K] int8, codes [N,K] in {-1,0,1} -> int64 [M,N]."""
    return q_x.astype(np.int64) @ codes.astype(np.int64).T


def fused_out(acc: np.ndarray, s_x: np.ndarray, scale, bias, clip_val: float) -> np.ndarray:
    """fp32 epilogue in the reference's order: *= scale, += bias, clip (binary/modules.py:105-108)."""
    y = acc.astype(np.float32) * s_x.astype(np.float32)[:, None]
    if scale is not None:
        y = y * np.asarray(scale, np.float32)[None, :]
    if bias is not None:
        y = y + np.asarray(bias, np.float32)[None, :]
    return np.clip(y, -np.float32(abs(clip_val)), np.float32(abs(clip_val))).astype(np.float32)


def codes_of(w: np.ndarray, mode: str = "ternary") -> np.ndarray:
    """Decision rules of binary/modules.py:5-26 on a numpy array (comparison in w's own dtype)."""
    w = np.asarray(w)
    if mode == "ternary":
        return (w >= 0.5).astype(np.int8) - (w <= -0.5).astype(np.int8)
    if mode in ("binary", "sign"):
        return ((w >= 0.0).astype(np.int8) * 2 - 1).astype(np.int8)
    if mode in ("binary01", "step"):
        return (w >= 0.0).astype(np.int8)
    raise ValueError(mode)
