"""Blocks of DEPENDENT quantised Linear layers: the reference wires the layer as
``fc1 -> gelu_new -> fc2`` (``PhiMLP``, impl/tf/moondream/layers.py:32-43) and ``Wqkv`` / ``out_proj``
(layers.py:181-186), and hands the activation function INTO the ternary layer, where it is applied after the clip
(impl/tf/moondream/ternary.py:245-246, 298-299; utils/conversion.py:185-196).

``ChainedLinears`` keeps that wiring and picks the launch form per call:
  * decode (M <= ``chain_max_tokens`` tokens, default 2, no autograd): ONE persistent launch for the whole chain
    (``tb200_chain_forward``): weights stream across the layer boundaries, activations are re-quantised to int8 inside
    the kernel, the activation function runs in the producing layer's epilogue;
  * anything else: layer by layer through the layers' own forward (GEMV / tcgen05 GEMM, or the fake-quant training
    formulation) with the activation as a torch op -- the same values.
There is no CPU path: the layers raise ``TernaryB200Error`` on CPU tensors.
"""
from __future__ import annotations

import torch
from torch import nn

from . import ops
from ._lib import TernaryB200Error
from .modules import _PackedQuantLinear, _require_cuda

_ACT_FNS = {
    None: lambda t: t,
    "none": lambda t: t,
    "gelu_new": lambda t: nn.functional.gelu(t, approximate="tanh"),
    "gelu_tanh": lambda t: nn.functional.gelu(t, approximate="tanh"),
    "gelu_pytorch_tanh": lambda t: nn.functional.gelu(t, approximate="tanh"),
    "relu": torch.relu,
}


class ChainedLinears(nn.Module):
    """``layers[i]`` are quantised Linear modules of this package with ``in_features[i+1] == out_features[i]``;
    ``activations[i]`` in {None, "gelu_new", "relu"} is applied to layer i's clipped output (ternary.py:245-246)."""

    def __init__(self, layers, activations=None):
        super().__init__()
        layers = list(layers)
        if not layers:
            raise ValueError("ChainedLinears needs at least one layer")
        activations = list(activations) if activations is not None else [None] * len(layers)
        if len(activations) != len(layers):
            raise ValueError("one activation entry per layer")
        for a in activations:
            if a not in _ACT_FNS:
                raise ValueError(f"unknown activation {a!r}; expected one of {sorted(k for k in _ACT_FNS if k)}")
        for i, l in enumerate(layers):
            if not isinstance(l, _PackedQuantLinear):
                raise TypeError(f"layer {i} is {type(l).__name__}, not a quantised Linear of ternary_vllm_b200")
            if i and l.in_features != layers[i - 1].out_features:
                raise ValueError(f"layer {i} consumes {l.in_features} features but layer {i - 1} produces "
                                 f"{layers[i - 1].out_features}")
        self.layers = nn.ModuleList(layers)
        self.activations = activations
        self._plans = {}            # (dtype, device) -> (key, ChainPlan)
        # every CTA of the chain kernel re-quantises all M activation rows of a layer for itself (~2 us per token and layer), so
        # above two tokens the layer-by-layer launches win (tools/chain_bench.py); the kernel itself serves up to CHAIN_MAX_M
        self.chain_max_tokens = 2

    # -- plan cache: rebuilt when any layer re-packs (the plan holds raw pointers into the packed buffers) --
    def _plan(self, dtype, device):
        packed = [l.packed_weights() for l in self.layers]

        def sig(t):
            return None if t is None else (t.data_ptr(), t._version, t.dtype)
        # the cached plan keeps the packed tensors it points to alive, so a re-pack always shows as a new address
        key = tuple((p.data_ptr(), sig(getattr(l, "scale", None)), sig(l.bias), l.maxval) for l, p in zip(self.layers, packed))
        hit = self._plans.get((dtype, device))
        if hit is not None and hit[0] == key:
            return hit[1]
        descs = []
        for l, p, a in zip(self.layers, packed, self.activations):
            if p.device != device:
                raise TernaryB200Error(f"packed weights live on {p.device} but the input is on {device}")
            descs.append((p, l.scale.detach() if l._has_scale else None, None if l.bias is None else l.bias.detach(), l.maxval, a))
        plan = ops.ChainPlan(descs, self.layers[0].in_features, dtype)
        self._plans[(dtype, device)] = (key, plan)
        return plan

    def _chain_ok(self, x) -> bool:
        if len(self.layers) > ops.L.CHAIN_MAX_LAYERS or x.dtype not in (torch.bfloat16, torch.float16, torch.float32):
            return False
        M = x.numel() // max(1, x.shape[-1])
        if M < 1 or M > self.chain_max_tokens:
            return False
        for l in self.layers:
            if l.pack_bits != 2 or not l.act_quant or (l._needs_autograd(x) and not l._frozen) or (l.training and not l._has_scale):
                return False
        return True

    def forward(self, x, pdl: bool = False):
        _require_cuda(x)
        if self._chain_ok(x):
            return self._plan(x.dtype, x.device).forward(x, pdl=pdl)
        for l, a in zip(self.layers, self.activations):
            x = _ACT_FNS[a](l(x))
        return x


class TernaryMLP(ChainedLinears):
    """``PhiMLP`` (impl/tf/moondream/layers.py:32-43): ``fc2(act(fc1(x)))`` with both linears quantised."""

    def __init__(self, fc1: _PackedQuantLinear, fc2: _PackedQuantLinear, hidden_act: str = "gelu_new"):
        super().__init__([fc1, fc2], [hidden_act, None])

    @property
    def fc1(self):
        return self.layers[0]

    @property
    def fc2(self):
        return self.layers[1]


# ---- wiring into an existing model (the HF Phi blocks the reference converts: quantize_moondream leaves `mlp.fc1` / `mlp.fc2`
# as quantised layers around `mlp.activation_fn`, modeling_phi.PhiMLP / impl/tf/moondream/layers.py:32-43) ----
_ACT_NAMES = {"NewGELUActivation": "gelu_new", "PytorchGELUTanh": "gelu_new", "GELUTanh": "gelu_new", "ReLU": "relu"}


def _activation_name(fn):
    """Name (key of ``_ACT_FNS``) of an activation module / callable this package can fuse, else None."""
    if fn is None:
        return None
    name = type(fn).__name__
    if name in _ACT_NAMES:
        return _ACT_NAMES[name]
    if isinstance(fn, nn.GELU):
        return "gelu_new" if getattr(fn, "approximate", "none") == "tanh" else None
    if fn is torch.relu or fn is nn.functional.relu:
        return "relu"
    return None


def fuse_mlp_blocks(model: nn.Module, max_tokens: int = 2, verbose: bool = False) -> int:
    """For every sub-module that looks like the reference's MLP block -- quantised children ``fc1`` and ``fc2`` with
    ``fc2.in_features == fc1.out_features`` and an ``activation_fn`` / ``act`` this package knows -- route decode-shaped,
    no-grad calls (at most ``max_tokens`` tokens) through ONE ``tb200_chain_forward`` launch; every other call keeps the
    module's own forward.  Call it after ``quantize_moondream`` / ``load_safetensors_packed``.  Returns the number of blocks
    fused.  The block's parameters and state dict are untouched (the fused path reads the same layers)."""
    n = 0
    for name, mod in list(model.named_modules()):
        fc1, fc2 = getattr(mod, "fc1", None), getattr(mod, "fc2", None)
        if not (isinstance(fc1, _PackedQuantLinear) and isinstance(fc2, _PackedQuantLinear)) or isinstance(mod, ChainedLinears):
            continue
        if fc2.in_features != fc1.out_features or getattr(mod, "_tb_fused_mlp", None) is not None:
            continue
        act = _activation_name(getattr(mod, "activation_fn", getattr(mod, "act", None)))
        if act is None:
            continue
        fused = TernaryMLP(fc1, fc2, act)
        fused.chain_max_tokens = max_tokens
        object.__setattr__(mod, "_tb_fused_mlp", fused)            # not a registered child: parameters stay where they are
        orig_forward = mod.forward

        def forward(x, *args, _fused=fused, _orig=orig_forward, **kwargs):
            if not args and not kwargs and torch.is_tensor(x) and x.is_cuda and _fused._chain_ok(x):
                return _fused(x)
            return _orig(x, *args, **kwargs)
        mod.forward = forward
        n += 1
        if verbose:
            print(f"fused {name}: fc1 -> {act} -> fc2 in one launch for <= {max_tokens} tokens")
    return n
