"""Model surgery and latent-weight clipping: drop-in for the reference's ``binary/utils.py``."""
from __future__ import annotations

from torch import nn

from . import ops
from .modules import (QUANT_CLASSES, ScaledBinary01Linear, ScaledBinaryLinear, ScaledTernaryLinear,
                      SmoothlyBinarizedLinear, linear_to_quantized)


def quantized_size_bits(phimodel: nn.Module) -> dict:
    """Parameter / bit accounting of quantize_moondream (binary/utils.py:36-58); ternary = 2 bits/weight."""
    acc = dict(n_bin_params=0, n_binary_modules=0, n_ternary_modules=0, n_linear_modules=0, n_bits=0)
    for mod in phimodel.modules():
        nb = 0 if getattr(mod, "bias", None) is None else mod.bias.numel()
        if isinstance(mod, (ScaledBinaryLinear, ScaledBinary01Linear)):
            acc["n_bin_params"] += mod.weights.numel()
            acc["n_binary_modules"] += 1
            acc["n_bits"] += mod.weights.numel() + 16 * 2 * nb
        elif isinstance(mod, SmoothlyBinarizedLinear):
            acc["n_bin_params"] += mod.weights.numel()
            acc["n_binary_modules"] += 1
            acc["n_bits"] += mod.weights.numel() + 16 * nb
        elif isinstance(mod, ScaledTernaryLinear):
            acc["n_ternary_modules"] += 1
            acc["n_bits"] += mod.weights.numel() * 2 + 16 * 2 * nb
        elif isinstance(mod, nn.Linear):
            acc["n_linear_modules"] += 1
            acc["n_bits"] += mod.weight.numel() * 16 + nb * 16
        elif isinstance(mod, (nn.Embedding, nn.LayerNorm)):
            acc["n_bits"] += sum(p.numel() for p in mod.parameters()) * 16
    return acc


def quantize_moondream(model, start_skip=1, last_skip=0, quantization="binary", scaling="none",
                       neuron_scale="uniform", remove_blocks=None, verbose=True):
    """Same contract as binary/utils.py:7-71: walk ``model.text_model``, skip ``lm_head`` and the
    first ``start_skip`` / last ``last_skip`` transformer blocks, convert every other direct
    ``nn.Linear`` child with :func:`linear_to_quantized`, optionally delete blocks and renumber
    ``mixer.layer_idx``, then print the size estimate.  Returns the accounting dict."""
    phimodel = model.text_model
    n_params_orig = sum(p.numel() for p in phimodel.parameters())
    n_blocks = len(model.text_model.transformer.h)
    layers_to_skip = [f"transformer.h.{i}." for i in range(start_skip)] + \
                     [f"transformer.h.{n_blocks - 1 - i}." for i in range(last_skip)]
    for name, mod in list(phimodel.named_modules()):
        if "lm_head" not in name and not any(skipped in name for skipped in layers_to_skip):
            if linear_to_quantized(mod, quantization=quantization, scaling=scaling, neuron_scale=neuron_scale) and verbose:
                print(f"Converted {name} to {quantization}")
    remove_blocks = remove_blocks or []
    phimodel.transformer.h = nn.ModuleList([m for i, m in enumerate(phimodel.transformer.h) if i not in remove_blocks])
    for i, mod in enumerate(phimodel.transformer.h):
        mod.mixer.layer_idx = i
    acc = quantized_size_bits(phimodel)
    acc["n_params"] = sum(p.numel() for p in phimodel.parameters())
    acc["estimated_model_size_gb"] = n_params_orig * 2 / 1024 ** 3
    acc["estimated_quantized_model_size_gb"] = acc["n_bits"] / 8 / 1024 ** 3
    if verbose:
        print(f"number of parameters: {acc['n_params']}")
        print(f"number of binary parameters: {acc['n_bin_params']}")
        print(f"number of binary modules: {acc['n_binary_modules']}")
        print(f"number of ternary modules: {acc['n_ternary_modules']}")
        print(f"number of linear module left: {acc['n_linear_modules']}")
        print(f"estimated original size: {acc['estimated_model_size_gb']:.3f} GB")
        print(f"estimated quantized size: {acc['estimated_quantized_model_size_gb']:.3f} GB")
    return acc


def clip_binary_weights(model: nn.Module):
    """binary/utils.py:74-77 — latent weights back into [-1, 1] after every optimiser step, done
    in place by one kernel per tensor; the packed cache is invalidated."""
    for mod in model.modules():
        if isinstance(mod, QUANT_CLASSES):
            w = mod.weights.data
            if not w.is_contiguous():
                w = w.contiguous()
                mod.weights.data = w
            ops.clip_latent_(w)
            mod.invalidate_packed()
