"""Numpy restatement of a chain of dependent quantised layers (fused MLP block).  TEST ORACLE ONLY.

The reference composes the layers as modules: ``fc2(act(fc1(x)))`` with the activation handed to the ternary
layer and applied after the clip (impl/tf/moondream/layers.py:32-43; impl/tf/moondream/ternary.py:245-246,298-299;
utils/conversion.py:185-196 passes ``keras.activations.gelu(approximate=True)``), every intermediate tensor in the
activation dtype (binary/modules.py:102-108 keeps ``x.dtype``).  With the per-token int8 activation quantisation
this repo defines (oracle/ref_int8.py), layer l is

    y_l   = fused_out( int_acc( quant_act(h_l), codes_l ), s_x, scale_l, bias_l, clip_l )      (fp32 epilogue)
    h_l+1 = round_dt( act_l( round_dt( y_l ) ) )           act = identity: a single rounding

gelu_new = 0.5 x (1 + tanh(sqrt(2/pi) (x + 0.044715 x^3))) evaluated in fp32 (layers.py:21-27).
"""
from __future__ import annotations

import numpy as np
import torch

from . import ref_int8

_DT = {"bf16": torch.bfloat16, "f16": torch.float16, "f32": torch.float32}


def round_dt(a: np.ndarray, dt: str) -> np.ndarray:
    if dt == "f32":
        return np.asarray(a, np.float32)
    return torch.from_numpy(np.asarray(a, np.float32)).to(_DT[dt]).float().numpy()


def gelu_tanh(x: np.ndarray) -> np.ndarray:
    x = np.asarray(x, np.float32)
    u = np.float32(0.7978845608028654) * (x + np.float32(0.044715) * x * x * x)
    return (np.float32(0.5) * x * (np.float32(1.0) + np.tanh(u, dtype=np.float32))).astype(np.float32)


def apply_act(y: np.ndarray, act, dt: str) -> np.ndarray:
    """y: fp32 epilogue values -> stored activation (already rounded to dt)."""
    if act in (None, "none"):
        return round_dt(y, dt)
    if act == "relu":
        return round_dt(np.maximum(y, np.float32(0.0)), dt)
    if act in ("gelu_new", "gelu_tanh"):
        return round_dt(gelu_tanh(round_dt(y, dt)), dt)
    raise ValueError(act)


def layer_forward(h: np.ndarray, codes: np.ndarray, scale, bias, clip, act, dt: str):
    """One layer on stored activations h [M,K] (fp32 view of dt values) -> (stored output, int64 accumulators)."""
    q, s, _ = ref_int8.quant_act(h)
    acc = ref_int8.int_acc(q, codes)
    y = ref_int8.fused_out(acc, s, scale, bias, clip)
    return apply_act(y, act, dt), acc


def chain_forward(x: np.ndarray, layers, dt: str = "bf16"):
    """layers: list of (codes [N,K] int8, scale|None, bias|None, clip, act).  Returns (out, [h_1 .. h_L], [acc_1 .. acc_L])."""
    h = np.asarray(x, np.float32)
    hs, accs = [], []
    for (codes, scale, bias, clip, act) in layers:
        h, acc = layer_forward(h, codes, scale, bias, clip, act, dt)
        hs.append(h)
        accs.append(acc)
    return h, hs, accs
