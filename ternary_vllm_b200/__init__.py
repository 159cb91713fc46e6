"""ternary_vllm_b200 — B200-native (sm_100a) quantised-Linear hot path of BenCrulis/ternary-vllm.

Public surface mirrors the reference's ``binary/modules.py`` and ``binary/utils.py`` plus the
operator set of its (absent) Larq-Compute-Engine fork; see INTEGRATION.md.
"""
from ._lib import TernaryB200Error, LIB_PATH  # noqa: F401
from . import ops  # noqa: F401
from .modules import (  # noqa: F401
    ScaledBinaryLinear, ScaledBinary01Linear, ScaledTernaryLinear, SmoothlyBinarizedLinear, SignSTESat,
    sign_ste, step_ste, tri_step_ste, sign_ste_sat, step_ste_sat,
    linear_to_quantized, compute_and_set_biobjective_gradient, compute_smooth_binary_stats,
)
from .checkpoint import load_safetensors_packed, export_packed  # noqa: F401
from .blocks import ChainedLinears, TernaryMLP, fuse_mlp_blocks  # noqa: F401
from .utils import quantize_moondream, clip_binary_weights, quantized_size_bits  # noqa: F401

__version__ = "0.1.0"
