"""Drop-in replacements for the reference's quantised Linear layers (``binary/modules.py``).

Same class names, constructor ``(in_features, out_features, bias=True, max_val=100.0)``,
``forward(x)`` and parameter names (``weights``, ``bias``, ``scale``, attribute ``maxval``)
as the reference (binary/modules.py:55-146), so ``isinstance`` checks, name-based checkpoint
loading (convert.py:75-79) and ``clip_binary_weights`` keep working.  What changes:

* inference (no autograd graph needed) runs on packed 2-bit weights through the C ABI:
  ``ops.ternary_matmul`` -> HBM-bound GEMV for <= 32 tokens, tcgen05 ``kind::i8`` GEMM above;
  the latent weights are quantised+packed once (one kernel) and re-packed only when they change;
* activations are quantised per token to int8 inside the kernels (the reference keeps them in
  floating point); ``act_quant=False`` selects the float-activation GEMV (decode shapes only);
* training (autograd enabled and a parameter or the input requires grad) keeps the
  reference's fake-quant straight-through formulation, evaluated with torch ops on the GPU;
* frozen weights work (the reference raises UnboundLocalError, modules.py:22-23).
"""
from __future__ import annotations

import torch
from torch import nn

from . import ops
from ._lib import TernaryB200Error

# --------------------------------------------------------------------------- STE quantisers
# Forward values and gradients of binary/modules.py:5-51 (the `t`-before-assignment bug of the
# `not x.requires_grad` branch is not reproduced: frozen inputs return the plain codes).


def _ste(q: torch.Tensor, x: torch.Tensor, x_c: torch.Tensor | None = None) -> torch.Tensor:
    if not x.requires_grad:
        return q
    x_c = x if x_c is None else x_c
    return q + x_c - x_c.detach()


def sign_ste(x):
    return _ste(((x >= 0.0) * 2.0 - 1.0).to(dtype=x.dtype), x)


def step_ste(x):
    return _ste((x >= 0.0).to(dtype=x.dtype), x)


def tri_step_ste(x):
    return _ste((x >= 0.5).to(dtype=x.dtype) - (x <= -0.5).to(dtype=x.dtype), x)


def sign_ste_sat(x):
    return _ste(((x >= 0.0) * 2.0 - 1.0).to(dtype=x.dtype), x, x.clip(-1.0, 1.0))


def step_ste_sat(x):
    return _ste((x >= 0.0).to(dtype=x.dtype), x, x.clip(-1.0, 1.0))


class SignSTESat(nn.Module):
    def forward(self, x):
        return sign_ste_sat(x)


def _require_cuda(x):
    if not x.is_cuda:
        raise TernaryB200Error("ternary_vllm_b200 layers run on CUDA (sm_100a) only; there is no CPU fallback")


# --------------------------------------------------------------------------- layers
class _Touch:
    """Write counter shared by a TrackedParameter and the layer that caches its packed form."""
    __slots__ = ("n",)

    def __init__(self):
        self.n = 0


class TrackedParameter(nn.Parameter):
    """``nn.Parameter`` whose ``.data`` accesses are observable.

    The reference mutates latent weights through ``.data`` (``p.data[:] = v`` when a checkpoint is loaded by
    parameter name, convert.py:76-78; ``weights.data = ...`` in linear_to_quantized and clip_binary_weights,
    binary/modules.py:226,261,292, binary/utils.py:77).  ``.data`` returns an alias with its OWN autograd version
    counter, so such writes never bump ``weights._version`` and a cached packed copy would silently go stale.
    Here every ``.data`` read or assignment bumps a counter that is part of the cache key, so the next forward
    re-packs (conservatively: a read-only ``.data`` access costs one re-pack).  In-place ops on the parameter itself
    (optimiser steps, ``load_state_dict``) are caught by ``_version`` as before.  Not covered: writing through an
    alias obtained BEFORE a forward and kept across it (``d = p.data; layer(x); d[:] = v``) -- call
    ``invalidate_packed()`` for that."""

    def __new__(cls, data=None, requires_grad=True):
        if data is None:
            data = torch.empty(0)
        p = torch.Tensor._make_subclass(cls, data, requires_grad)
        p._tb_touch = _Touch()
        return p

    @property
    def data(self):
        self._tb_touch.n += 1
        return torch.Tensor.data.__get__(self)

    @data.setter
    def data(self, value):
        self._tb_touch.n += 1
        torch.Tensor.data.__set__(self, value)

    def __deepcopy__(self, memo):
        r = type(self)(torch.Tensor.data.__get__(self).clone(memory_format=torch.preserve_format), self.requires_grad)
        memo[id(self)] = r
        return r

    def __reduce_ex__(self, proto):
        return (TrackedParameter, (torch.Tensor.data.__get__(self), self.requires_grad))


class _PackedQuantLinear(nn.Module):
    """Shared machinery: latent fp weights + cached 2-bit packed form + fused kernels."""

    _mode = "ternary"          # key of ops._MODES
    _ste = staticmethod(tri_step_ste)
    _has_scale = True

    def __init__(self, in_features: int, out_features: int, bias: bool = True, max_val=100.0):
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.weights = TrackedParameter(torch.randn(out_features, in_features))
        self.bias = nn.Parameter(torch.randn(out_features)) if bias else None
        if self._has_scale:
            self.scale = nn.Parameter(torch.full((out_features,), 1.0))
        self.maxval = abs(max_val)
        self.act_quant = True
        # bits per packed weight: 2 (format C, every kernel) or -- two-valued layers only -- 1 (format B1: half the HBM bytes
        # of a decode GEMV; binary/utils.py:42-45 accounts these layers at one bit).  Change it with set_pack_bits().
        self.pack_bits = 2
        # the packed weights are a BUFFER: they follow .to()/.cuda() with the module; not part of the state dict (the
        # reference's checkpoint keys are weights / bias / scale only) until freeze() makes them the deployed form
        self.register_buffer("_packed", None, persistent=False)
        self._packed_key = None
        self._frozen = False
        self._released = False

    # -- packed-weight cache ------------------------------------------------------------
    def _key(self):
        w = self.weights
        try:
            version = w._version
        except RuntimeError:           # parameters created under torch.inference_mode() have no version counter
            version = -1
        touch = getattr(w, "_tb_touch", None)
        return (w.data_ptr() if w.numel() else 0, version, touch.n if touch is not None else -1,
                tuple(w.shape), w.dtype, w.device)

    def invalidate_packed(self):
        """Force a re-pack on the next forward.  Needed only for writes this module cannot see (see
        TrackedParameter); ``.data`` writes, in-place ops, ``load_state_dict`` and ``.to()`` are tracked."""
        if self._released:
            raise TernaryB200Error("the latent weights of this layer were released by freeze(release_latent=True); "
                                   "it can no longer be re-packed")
        self._packed_key = None
        self._frozen = False
        if "_packed" in self._buffers:
            self._non_persistent_buffers_set.add("_packed")

    def set_pack_bits(self, bits: int):
        """1: store the (two-valued) weights at one bit each (format B1, ``tb200_gemv_b1``); 2: the default format C."""
        if bits not in (1, 2) or (bits == 1 and self._mode == "ternary"):
            raise ValueError("pack_bits must be 2, or 1 for a binary / binary01 layer")
        if bits != self.pack_bits:
            if self._released:
                raise TernaryB200Error("latent weights were released; the layer cannot be re-packed at another width")
            self.pack_bits = bits
            self._packed_key = None
            self._packed = None
        return self

    def packed_weights(self) -> torch.Tensor:
        """uint8 ``[out_features, pitch]`` in the reference's format C (or B1 when ``pack_bits == 1``); rebuilt when stale."""
        if self._frozen and self._packed is not None:
            return self._packed
        key = self._key() + (self.pack_bits,)
        if self._packed is None or self._packed_key != key:
            if self._released:
                raise TernaryB200Error("latent weights were released; the packed weights are missing")
            w = torch.Tensor.data.__get__(self.weights)
            self._packed = ops.quantize_pack_b1(w) if self.pack_bits == 1 else ops.quantize_pack(w, self._mode)
            self._packed_key = key
        return self._packed

    def freeze(self, release_latent: bool = False):
        """Pack now and stop checking for weight updates (deployment).  The packed weights become a persistent
        buffer (saved by ``state_dict()`` as ``_packed``, restored by ``load_state_dict``).  ``release_latent``
        additionally drops the latent fp weights (the module can then no longer train or re-pack)."""
        self.packed_weights()
        self._frozen = True
        self._non_persistent_buffers_set.discard("_packed")
        if release_latent:
            self.weights = nn.Parameter(torch.empty(0, device=self._packed.device), requires_grad=False)
            self._released = True
        return self

    def _load_from_state_dict(self, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs):
        pk = state_dict.get(prefix + "_packed")
        if pk is not None and not self._frozen:
            # a checkpoint of a frozen layer: adopt its deployed form
            self._packed = torch.empty_like(pk, device=self.weights.device if self._packed is None else self._packed.device)
            self._frozen = True
            self._non_persistent_buffers_set.discard("_packed")
            # the checkpoint does not name the packing width: the row pitch tells (formats B1 and C differ for K > 64)
            p1, p2 = ops.b1_pitch(self.in_features), ops.packed_pitch(self.in_features)
            if pk.shape[-1] == p1 and p1 != p2 and self._mode != "ternary":
                self.pack_bits = 1
            elif pk.shape[-1] == p2 and p1 != p2:
                self.pack_bits = 2
            w = state_dict.get(prefix + "weights")
            if w is not None and w.numel() == 0:
                self.weights = nn.Parameter(torch.empty(0, device=self._packed.device), requires_grad=False)
                self._released = True
        super()._load_from_state_dict(state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs)

    def codes(self) -> torch.Tensor:
        """int8 ``[out_features, in_features]`` quantised weights (ops.unpack_ternary)."""
        pw = self.packed_weights()
        if self.pack_bits == 1:
            pw = ops.expand_b1(pw, self.in_features, self._mode == "binary01")
        return ops.unpack(pw, self.in_features)

    # -- forward ---------------------------------------------------------------------------
    def _needs_autograd(self, x) -> bool:
        if not torch.is_grad_enabled():
            return False
        ps = [self.weights, self.bias, getattr(self, "scale", None)]
        return x.requires_grad or any(p is not None and p.requires_grad for p in ps)

    def _forward_ste(self, x):
        # the reference's fake-quant formulation (binary/modules.py:63-71, 82-90, 101-109)
        w = self._ste(self.weights).to(x.dtype)
        y = nn.functional.linear(x, w)
        if self._has_scale:
            y = y * self.scale.to(y.dtype)
        if self.bias is not None:
            y = y + self.bias.to(y.dtype)
        return y.clip(-self.maxval, self.maxval)

    def _forward_packed(self, x):
        if not x.is_cuda:
            raise TernaryB200Error("ternary_vllm_b200 layers run on CUDA (sm_100a) only; there is no CPU fallback")
        scale = self.scale if self._has_scale else None
        pw = self.packed_weights()
        if pw.device != x.device:
            raise TernaryB200Error(f"packed weights live on {pw.device} but the input is on {x.device}; move the "
                                   "layer with .to() (the packed buffer follows the module)")
        if self.pack_bits == 1:
            if not self.act_quant:
                pw = ops.expand_b1(pw, self.in_features, self._mode == "binary01")
            else:
                return ops.binary_matmul(x, pw, self.in_features, scale, self.bias, self.maxval, self._mode == "binary01")
        if not self.act_quant:
            x2 = x.reshape(-1, x.shape[-1])
            if x2.shape[0] > ops.L.GEMV_MAX_M:
                raise TernaryB200Error("act_quant=False supports at most 32 tokens per call")
            y = ops.gemv_f32(x2.contiguous(), pw, scale, self.bias, self.maxval)
            return y.reshape(*x.shape[:-1], self.out_features)
        return ops.ternary_matmul(x, pw, scale, self.bias, self.maxval)

    def forward(self, x):
        _require_cuda(x)
        if self._needs_autograd(x) and not self._frozen:
            return self._forward_ste(x)
        return self._forward_packed(x)

    def extra_repr(self):
        return f"in_features={self.in_features}, out_features={self.out_features}, bias={self.bias is not None}, maxval={self.maxval}"


class ScaledBinaryLinear(_PackedQuantLinear):
    """binary/modules.py:55-71 — weights in {-1,+1} (sign, -0.0 -> +1), per-channel scale."""
    _mode = "binary"
    _ste = staticmethod(sign_ste)


class ScaledBinary01Linear(_PackedQuantLinear):
    """binary/modules.py:74-90 — weights in {0,+1}."""
    _mode = "binary01"
    _ste = staticmethod(step_ste)


class ScaledTernaryLinear(_PackedQuantLinear):
    """binary/modules.py:93-109 — weights in {-1,0,+1} with thresholds +-0.5."""
    _mode = "ternary"
    _ste = staticmethod(tri_step_ste)


class SmoothlyBinarizedLinear(_PackedQuantLinear):
    """binary/modules.py:112-146 — no scale; train: clip(W,+-1) dense, eval: sign(W)."""
    _mode = "binary"
    _has_scale = False

    def count_converged_parameters(self, eps):
        return (1.0 - self.weights.abs() < eps).sum()

    def count_binary_parameters(self):
        return self.weights.numel()

    def compute_binarization_gradient(self):
        return -torch.sign(self.weights)

    def compute_and_set_biobjective_gradient(self):
        # deprecated in the reference too (modules.py:128-137)
        g = self.weights.grad
        bg = self.compute_binarization_gradient()
        g_norm = torch.square(g).sum().sqrt()
        ng = g / g_norm if g_norm > 0.0 else g * 0.0
        bg_norm = torch.square(bg).sum().sqrt()
        nbg = bg / bg_norm if bg_norm > 0.0 else bg * 0.0
        self.weights.grad = (ng + nbg) * (g_norm + bg_norm) * 0.5

    def forward(self, x):
        _require_cuda(x)
        if self.training:
            w = self.weights.clip(-1.0, 1.0).to(x.dtype)
            y = nn.functional.linear(x, w, bias=None if self.bias is None else self.bias.to(x.dtype))
            return y.clip(-self.maxval, self.maxval)
        if self._needs_autograd(x) and not self._frozen:
            w = (self.weights >= 0.0).to(x.dtype) * 2.0 - 1.0
            y = nn.functional.linear(x, w, bias=None if self.bias is None else self.bias.to(x.dtype))
            return y.clip(-self.maxval, self.maxval)
        return self._forward_packed(x)


QUANT_CLASSES = (ScaledBinaryLinear, ScaledBinary01Linear, ScaledTernaryLinear, SmoothlyBinarizedLinear)


# --------------------------------------------------------------------------- training helpers
def _biobjective_fused(model: nn.Module) -> bool:
    """GPU form of compute_and_set_biobjective_gradient: one reduction launch per gradient / smooth-binary weight tensor
    into two device accumulators, then one in-place launch per gradient (``tb200_sumsq_accumulate`` /
    ``tb200_biobjective_apply``); no host synchronisation.  Returns False (nothing touched) unless every tensor involved is
    a contiguous CUDA tensor of a supported dtype on one device with the standard ``-sign(weights)`` binarisation gradient."""
    grads, smooth = [], []
    dev = None
    for mod in model.modules():
        is_smooth = hasattr(mod, "compute_binarization_gradient")
        if is_smooth and type(mod).compute_binarization_gradient is not SmoothlyBinarizedLinear.compute_binarization_gradient:
            return False
        for name, p in mod.named_parameters(recurse=False):
            if is_smooth and name == "weights":
                if p.grad is None:
                    return False
                smooth.append(p)
            if p.grad is not None:
                grads.append((p, is_smooth and name == "weights"))
        if is_smooth and not any(n == "weights" for n, _ in mod.named_parameters(recurse=False)):
            return False
    if not grads:
        return False
    for p, _ in grads:
        g = p.grad
        if not (g.is_cuda and g.is_contiguous() and p.is_contiguous() and g.dtype == p.dtype
                and g.dtype in (torch.float32, torch.bfloat16, torch.float16)):
            return False
        if dev is None:
            dev = g.device
        if g.device != dev or p.device != dev:
            return False
    acc = torch.zeros(2, dtype=torch.float64, device=dev)
    ng_sq, nbg_sq = acc[0], acc[1]
    for p, _ in grads:
        ops.sumsq_accumulate_(ng_sq, p.grad)
    for p in smooth:
        ops.sumsq_accumulate_(nbg_sq, torch.Tensor.data.__get__(p) if isinstance(p, TrackedParameter) else p.data, count_nonzero=True)
    for p, is_w in grads:
        w = (torch.Tensor.data.__get__(p) if isinstance(p, TrackedParameter) else p.data) if is_w else None
        ops.biobjective_apply_(p.grad, w, ng_sq, nbg_sq)
    return True


@torch.no_grad()
def compute_and_set_biobjective_gradient(model: nn.Module):
    """binary/modules.py:149-176: blend the task gradient with the binarisation gradient
    ``-sign(w)`` at equal global norm.  CUDA gradients take the fused kernels (``_biobjective_fused``); anything else the
    reference's formulation in torch ops below."""
    if _biobjective_fused(model):
        return
    norm_g_sq = 0.0
    norm_bg_sq = 0.0
    bgs = {}
    for mod in model.modules():
        if hasattr(mod, "compute_binarization_gradient"):
            bg = mod.compute_binarization_gradient()
            bgs[mod] = bg
            norm_bg_sq = norm_bg_sq + torch.square(bg).sum()
        for p in mod.parameters(recurse=False):
            if p.grad is not None:
                norm_g_sq = norm_g_sq + torch.square(p.grad).sum()
    norm_g = norm_g_sq.sqrt()
    norm_bg = norm_bg_sq.sqrt()
    for mod in model.modules():
        if hasattr(mod, "compute_binarization_gradient"):
            bg = bgs[mod]
            g = mod.weights.grad
            ng = g / norm_g if norm_g > 0.0 else g * 0.0
            nbg = bg / norm_bg if norm_bg > 0.0 else bg * 0.0
            mod.weights.grad = (ng + nbg) * (norm_g + norm_bg) * 0.5
        for p in mod.parameters(recurse=False):
            if p.grad is not None:
                p.grad = (p.grad / norm_g) * (norm_g + norm_bg) if norm_g > 0.0 else p.grad


def compute_smooth_binary_stats(model, eps=1e-3):
    """binary/modules.py:179-186"""
    n_smooth_binary_params = 0
    n_converged = 0
    for mod in model.modules():
        if hasattr(mod, "count_converged_parameters"):
            n_smooth_binary_params += mod.count_binary_parameters()
            n_converged += mod.count_converged_parameters(eps)
    return n_smooth_binary_params, n_converged


# --------------------------------------------------------------------------- model surgery
def _prescale(weight, scaling):
    if scaling is None or scaling == "none":
        return 1.0, False
    if isinstance(scaling, float):
        return scaling, False
    if scaling == "std":
        return float(1.0 / weight.data.std()), False
    if scaling == "natural":
        return 1.0, True
    raise ValueError(f"Unknown scaling method: {scaling}")


def linear_to_quantized(model: nn.Module, quantization="binary", scaling="none", neuron_scale="uniform",
                        kmeans_iter=10):
    """Replace the direct ``nn.Linear`` children of ``model`` (binary/modules.py:189-299).
    Scale initialisation and latent-weight clipping run in the weight-prep kernels
    (``tb200_ternary_init`` / ``tb200_binary_init``).  Returns whether anything was converted."""
    converted = False
    for name, mod in model.named_children():
        if not isinstance(mod, nn.Linear):
            continue
        has_bias = mod.bias is not None
        dev = mod.weight.device
        if quantization in ("binary", "binary01"):
            if neuron_scale not in ("uniform", "independent"):
                raise ValueError(f"Unknown neuron scaling method: {neuron_scale}")
            pre, natural = _prescale(mod.weight, scaling)
            cls = ScaledBinaryLinear if quantization == "binary" else ScaledBinary01Linear
            new = cls(mod.in_features, mod.out_features, has_bias).to(dev)
            latent, scale = ops.binary_init(mod.weight, per_row=neuron_scale == "independent",
                                            positive_only=quantization == "binary01", prescale=pre, natural=natural)
            new.weights.data = latent
            new.scale.data = scale
        elif quantization == "smoothBinary":
            new = SmoothlyBinarizedLinear(mod.in_features, mod.out_features, has_bias).to(dev)
            new.weights.data = mod.weight.detach().to(torch.float32).clone()
        elif quantization == "ternary":
            new = ScaledTernaryLinear(mod.in_features, mod.out_features, has_bias).to(dev)
            latent, scale = ops.ternary_init(mod.weight, per_row=neuron_scale != "uniform", iters=kmeans_iter)
            new.weights.data = latent
            new.scale.data = scale
        else:
            continue   # the reference silently ignores unknown quantisation names
        if has_bias:
            new.bias.data = mod.bias.data
        new.invalidate_packed()
        model.add_module(name, new)
        converted = True
    return converted
