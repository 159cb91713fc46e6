"""Checkpoint path of the quantised layers.

``load_safetensors_packed`` is the reference's load-by-parameter-name loop (convert.py:72-79:
``for name, p in model.named_parameters(): p.data[:] = md_weights.get_tensor(name)``) with one change: the latent weights
of a quantised Linear never become resident.  Each ``...weights`` tensor is read from the file, moved to the GPU, turned
into 2-bit codes by ``tb200_quantize_pack`` straight into the layer's packed buffer and dropped again, so a 7B-class
checkpoint needs its packed size (2 bits per weight) plus ONE layer of latent weights in HBM instead of 4 bytes per weight.
Every other parameter is copied as the reference does.

``export_packed`` emits what ``utils/conversion.py:44-94`` feeds to the Keras layers: per quantised layer ``scale``,
``bias`` and the packed weights in the reference's formats A / B (planar over the flattened tensor,
impl/tf/moondream/ternary.py:57-175) or C (per row, ternary.py:374-398), produced by ``tb200_pack_format`` on the GPU.
"""
from __future__ import annotations

import torch
from torch import nn

from . import ops
from ._lib import TernaryB200Error
from .modules import QUANT_CLASSES


def _quant_layers(model: nn.Module) -> dict:
    return {name: m for name, m in model.named_modules() if isinstance(m, QUANT_CLASSES)}


@torch.no_grad()
def load_safetensors_packed(model: nn.Module, path, device=None, release_latent: bool = True, strict: bool = True) -> dict:
    """Load ``path`` (a .safetensors file) into ``model`` by parameter name.  Quantised layers end up FROZEN on their
    packed weights (``release_latent=True`` also drops the latent tensor, as ``freeze(release_latent=True)`` does);
    returns ``{"packed_layers", "copied", "missing", "bytes_packed", "bytes_latent_peak"}``.  ``strict`` raises on a
    parameter that the file lacks (the reference's ``get_tensor`` raises too)."""
    import safetensors

    layers = _quant_layers(model)
    weight_owner = {(name + ".weights" if name else "weights"): m for name, m in layers.items()}
    stats = {"packed_layers": 0, "copied": 0, "missing": [], "bytes_packed": 0, "bytes_latent_peak": 0}
    with safetensors.safe_open(str(path), "pt") as f:
        keys = set(f.keys())
        for name, p in list(model.named_parameters()):
            if name not in keys:
                if strict:
                    raise KeyError(f"{name} is not in {path}")
                stats["missing"].append(name)
                continue
            owner = weight_owner.get(name)
            if owner is None:
                p.data[:] = f.get_tensor(name).to(p.device, p.dtype)           # convert.py:78
                stats["copied"] += 1
                continue
            dev = torch.device(device) if device is not None else p.device
            if dev.type != "cuda":
                raise TernaryB200Error("packed loading needs a CUDA device for the layer (there is no CPU packer); "
                                       "move the model to the GPU first or pass device=")
            v = f.get_tensor(name)
            if tuple(v.shape) != (owner.out_features, owner.in_features):
                raise ValueError(f"{name}: checkpoint shape {tuple(v.shape)} != layer shape {(owner.out_features, owner.in_features)}")
            w = v.to(dev, non_blocking=True)
            stats["bytes_latent_peak"] = max(stats["bytes_latent_peak"], w.numel() * w.element_size())
            # the only pass over the latent weights
            owner._packed = ops.quantize_pack_b1(w.contiguous()) if owner.pack_bits == 1 else ops.quantize_pack(w.contiguous(), owner._mode)
            owner._packed_key = None
            owner._frozen = True
            owner._non_persistent_buffers_set.discard("_packed")
            if release_latent:
                owner.weights = nn.Parameter(torch.empty(0, device=dev), requires_grad=False)
                owner._released = True
            else:
                p.data[:] = w.to(p.dtype)
                owner._packed_key = owner._key() + (owner.pack_bits,)
            del w, v
            stats["packed_layers"] += 1
            stats["bytes_packed"] += owner._packed.numel()
    return stats


@torch.no_grad()
def export_packed(model: nn.Module, fmt: str = "C") -> dict:
    """``{layer name: {"scale", "bias", "packed", "clip_val", "shape"}}`` with ``packed`` a uint8 CPU tensor in the
    reference's format ``fmt`` ("A", "B": planar over the flattened [N*K] codes; "C": [N, ceil(K/4)] per row) --
    the three tensors ``dense_keras.set_weights([scale, b, packed_weights])`` takes (utils/conversion.py:44-94)."""
    out = {}
    for name, m in _quant_layers(model).items():
        codes = m.codes()                                                       # int8 [N, K] on the GPU
        packed = ops.pack_format(codes, fmt)
        out[name] = {"scale": None if not m._has_scale else m.scale.detach().float().cpu(),
                     "bias": None if m.bias is None else m.bias.detach().float().cpu(),
                     "packed": packed.cpu(), "clip_val": float(m.maxval), "shape": (m.out_features, m.in_features)}
    return out
