"""PyTorch-CPU restatement of the reference's fake-quant Linear path.  TEST ORACLE ONLY.

Restates ``/root/reference/binary/modules.py`` and ``binary/utils.py`` as pure
functions (no nn.Module surgery) so tests can compare the CUDA path with them on
the same seeded inputs.  PINNED: ``tests/test_oracle_vs_reference.py`` checks
every function here against the reference imported live (build container), and
``tests/golden/*.npz`` hold outputs of the live reference for the GPU box.

Conventions: weights are ``[N, K]`` (= ``[out_features, in_features]``), exactly
as the reference stores them.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

# --------------------------------------------------------------------------
# quantisers (binary/modules.py:5-51), forward values only
# --------------------------------------------------------------------------


def ternary_codes(w: torch.Tensor) -> torch.Tensor:
    """Decision rule of ``tri_step_ste`` (modules.py:21-26): +1 if w >= 0.5, -1 if
    w <= -0.5, else 0.  Returned as int8; this is what "bit-exact" refers to."""
    return (w >= 0.5).to(torch.int8) - (w <= -0.5).to(torch.int8)


def sign_codes(w: torch.Tensor) -> torch.Tensor:
    """``sign_ste`` (modules.py:5-10): +1 if w >= 0 (so -0.0 -> +1), else -1."""
    return (w >= 0.0).to(torch.int8) * 2 - 1


def step_codes(w: torch.Tensor) -> torch.Tensor:
    """``step_ste`` (modules.py:13-18): 1 if w >= 0 else 0."""
    return (w >= 0.0).to(torch.int8)


def ste_float(w: torch.Tensor, codes: torch.Tensor) -> torch.Tensor:
    """The float tensor the reference actually multiplies with: ``q + w - w.detach()``
    evaluated left to right in w.dtype (modules.py:10,18,26).  Differs from ``q`` by
    rounding noise (<= 1 ulp of |w|+1); kept so the oracle can reproduce the
    reference output to the last bit when needed."""
    q = codes.to(w.dtype)
    return (q + w) - w


def ste_float_sat(w: torch.Tensor, codes: torch.Tensor) -> torch.Tensor:
    """``*_ste_sat`` / ``SignSTESat`` (modules.py:29-51): gradient path clipped to
    [-1, 1]; forward value is ``q + clip(w) - clip(w)``."""
    c = w.clip(-1.0, 1.0)
    return (codes.to(w.dtype) + c) - c


# --------------------------------------------------------------------------
# layer forwards (binary/modules.py:55-146)
# --------------------------------------------------------------------------


def scaled_linear_forward(x, weights, scale, bias, maxval=100.0, kind="ternary", exact_ste=True):
    """Forward of Scaled{Ternary,Binary,Binary01}Linear (modules.py:101-109, 63-71,
    82-90).  ``exact_ste=True`` multiplies with ``q + w - w`` like the reference;
    ``False`` multiplies with the exact codes."""
    codes = {"ternary": ternary_codes, "binary": sign_codes, "binary01": step_codes}[kind](weights)
    wq = ste_float(weights, codes) if exact_ste else codes.to(weights.dtype)
    y = F.linear(x, wq.to(x.dtype))
    # in-place on the fresh F.linear output, like the reference: the result keeps
    # x.dtype even when scale/bias are fp32 (modules.py:105-107)
    y *= scale
    if bias is not None:
        y += bias
    return y.clip(-abs(maxval), abs(maxval))


def smooth_binary_forward(x, weights, bias, maxval=100.0, training=False):
    """SmoothlyBinarizedLinear.forward (modules.py:139-146): no scale."""
    if training:
        w = weights.clip(-1.0, 1.0)
    else:
        w = (weights >= 0.0).to(x.dtype) * 2.0 - 1.0
    y = F.linear(x, w, bias=bias)
    return y.clip(-abs(maxval), abs(maxval))


def packed_ternary_matmul(x, codes, scale, bias, clip_val):
    """Contract of ``ops.ternary_matmul(x, pw, scale, bias, clip_val)`` as spelled out by
    ScaledTernaryTf.call (ternary.py:234-244) and the commented fallback
    (ternary.py:289-297): clip((x @ unpack(pw).T) * scale + bias, +-clip_val)."""
    y = x @ codes.to(x.dtype).T
    y = y * scale + bias
    return y.clip(-clip_val, clip_val)


def ternary_mm(m: torch.Tensor, tm: torch.Tensor) -> torch.Tensor:
    """``ternary_mm`` of matrix_mul_tests.py:29-37: signed column sums, one output
    column per row of ``tm``.  (The reference allocates ``tm.shape[1]`` output columns
    and loops ``tm.shape[0]`` times; it is only ever called on square matrices.)"""
    res = torch.zeros(m.shape[0], tm.shape[0])
    pos, neg = tm > 0.0, tm < 0.0
    for i in range(tm.shape[0]):
        res[:, i] = m[:, pos[i]].sum(dim=1) - m[:, neg[i]].sum(dim=1)
    return res


# --------------------------------------------------------------------------
# weight preparation (binary/modules.py:189-299)
# --------------------------------------------------------------------------


def ternary_init(weight: torch.Tensor, neuron_scale="uniform", kmeans_iter=10):
    """Ternary branch of ``linear_to_quantized`` (modules.py:275-298).

    Returns ``(latent [N,K], scale)`` where ``scale`` is what the reference writes into
    ``st.scale.data[:]`` (a length-N vector after broadcast).  Any ``neuron_scale``
    other than "uniform" means per-row (the reference only tests for "uniform")."""
    a = weight.abs().detach()
    if neuron_scale == "uniform":
        a = a.flatten().unsqueeze(0)
    m = a.mean(dim=1)
    for _ in range(kmeans_iter):
        mask = (a - m[:, None]) > -m[:, None] / 2.0
        m = (a * mask).float().sum(dim=1) / mask.sum(dim=1)
    latent = (weight * (1.0 / m)[:, None]).clip(-1.0, 1.0).detach()
    # the freshly built ScaledTernaryLinear holds fp32 parameters (modules.py:277): the
    # assignments st.weights.data[:] = ..., st.scale.data[:] = m cast into fp32
    scale = torch.empty(weight.shape[0], dtype=torch.float32)
    scale[:] = m
    return latent.to(torch.float32), scale


def _prescale(weight, scaling, scale_vec):
    # modules.py:214-224 / 249-259
    if scaling is None or scaling == "none":
        return 1.0
    if isinstance(scaling, float):
        return scaling
    if scaling == "std":
        return 1.0 / weight.data.std()
    if scaling == "natural":
        return 1.0 / scale_vec[:, None]
    raise ValueError(f"Unknown scaling method: {scaling}")


def binary_init(weight: torch.Tensor, neuron_scale="uniform", scaling="none", kind="binary"):
    """Binary (modules.py:193-231) and binary01 (:232-266) branches.  Returns
    ``(latent, scale[N])``."""
    N, K = weight.shape
    pos = weight > 0
    neg = weight < 0
    if neuron_scale == "uniform":
        m_pos = (pos * weight).float().sum()
        m_neg = (neg * weight).float().sum()
        denom = K * N
    elif neuron_scale == "independent":
        m_pos = (pos * weight).float().sum(dim=1)
        m_neg = (neg * weight).float().sum(dim=1)
        denom = K
    else:
        raise ValueError(f"Unknown neuron scaling method: {neuron_scale}")
    scale = torch.empty(N, dtype=torch.float32)      # sb.scale is an fp32 parameter (modules.py:194)
    if kind == "binary":
        scale[:] = (m_pos - m_neg) / denom
    elif kind == "binary01":
        scale[:] = m_pos / denom
    else:
        raise ValueError(kind)
    s = _prescale(weight, scaling, scale)
    latent = (weight * s).clip(-1.0, 1.0).detach()
    return latent.to(torch.float32), scale


def clip_latent(w: torch.Tensor) -> torch.Tensor:
    """``clip_binary_weights`` per tensor (binary/utils.py:74-77)."""
    return w.clip(-1.0, 1.0)


# --------------------------------------------------------------------------
# training-side helpers (binary/modules.py:119-137, 149-186)
# --------------------------------------------------------------------------


def binarization_gradient(w):
    """modules.py:125-126"""
    return -torch.sign(w)


def count_converged(w, eps):
    """modules.py:119-120"""
    return (1.0 - w.abs() < eps).sum()


def biobjective_gradients(smooth_weights, smooth_grads, other_grads):
    """``compute_and_set_biobjective_gradient(model)`` (modules.py:149-176) on explicit
    lists: ``smooth_weights[i]``/``smooth_grads[i]`` belong to SmoothlyBinarizedLinear
    modules, ``other_grads`` are all remaining ``.grad`` tensors (including the biases of
    the smooth modules).  Returns (new_smooth_grads, new_other_grads).

    Note the reference first overwrites ``mod.weights.grad`` with the blend and then,
    in the same loop body, rescales *every* direct parameter grad of that module —
    including the blended one (modules.py:172-176)."""
    ng2 = 0.0
    nb2 = 0.0
    bgs = [binarization_gradient(w) for w in smooth_weights]
    for bg in bgs:
        nb2 = nb2 + torch.square(bg).sum()
    for g in list(smooth_grads) + list(other_grads):
        ng2 = ng2 + torch.square(g).sum()
    ng = ng2.sqrt() if torch.is_tensor(ng2) else torch.tensor(ng2).sqrt()
    nb = nb2.sqrt() if torch.is_tensor(nb2) else torch.tensor(nb2).sqrt()

    def rescale(g):
        return (g / ng) * (ng + nb) if ng > 0.0 else g

    out_s = []
    for g, bg in zip(smooth_grads, bgs):
        a = g / ng if ng > 0.0 else g * 0.0
        b = bg / nb if nb > 0.0 else bg * 0.0
        out_s.append(rescale((a + b) * (ng + nb) * 0.5))
    out_o = [rescale(g) for g in other_grads]
    return out_s, out_o


# --------------------------------------------------------------------------
# size accounting (binary/utils.py:36-58), per module
# --------------------------------------------------------------------------


def module_bits(kind: str, n_weights: int, n_bias: int) -> int:
    """Bit accounting of ``quantize_moondream`` (utils.py:42-56)."""
    if kind in ("binary", "binary01"):
        return n_weights + 16 * 2 * n_bias
    if kind == "smoothBinary":
        return n_weights + 16 * n_bias
    if kind == "ternary":
        return n_weights * 2 + 16 * 2 * n_bias
    if kind == "linear":
        return n_weights * 16 + n_bias * 16
    raise ValueError(kind)
