"""Thin tensor-level wrappers over the C ABI (``include/ternary_b200.h``).

PyTorch is plumbing here: it owns device memory and the current stream; every function
passes raw device pointers to ``libternary_b200.so``.  Nothing in this module computes on
the CPU and nothing falls back: non-CUDA tensors raise ``TernaryB200Error``.

The op set mirrors the reference's operator boundary:
``ternary_matmul`` = ``larq_compute_engine.ops.ternary_matmul`` (impl/tf/moondream/ternary.py:281),
``unpack`` = ``ops.unpack_ternary`` (:334), ``quantize_pack`` = ternarise + ``ops.pack_fn``
(utils/conversion.py:52-53,71-72).
"""
from __future__ import annotations

import torch

from . import _lib as L
from ._lib import TernaryB200Error

_DT = {torch.float32: L.TB200_F32, torch.bfloat16: L.TB200_BF16, torch.float16: L.TB200_F16}
_MODES = {"ternary": L.Q_TERNARY, "binary": L.Q_SIGN, "sign": L.Q_SIGN, "binary01": L.Q_STEP, "step": L.Q_STEP}
_FMTS = {"A": L.FMT_A, "B": L.FMT_B, "C": L.FMT_C}


def _dt(t: torch.Tensor) -> int:
    try:
        return _DT[t.dtype]
    except KeyError:
        raise TernaryB200Error(f"unsupported dtype {t.dtype}; use float32, bfloat16 or float16") from None


def _cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise TernaryB200Error("ternary_vllm_b200 runs on CUDA (sm_100a) tensors only; there is no CPU fallback")


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _ptr(t):
    return None if t is None else t.data_ptr()


def packed_pitch(K: int) -> int:
    """Row pitch (bytes) of the internal packed layout: ceil(K/4) rounded up to 16 so that every
    row starts on a 128-bit boundary (padding bytes hold code 1 = weight 0)."""
    return ((K + 3) // 4 + 15) // 16 * 16


# ------------------------------------------------------------------ weight preparation
def quantize_pack(w: torch.Tensor, mode: str = "ternary", ld_packed: int | None = None,
                  return_codes: bool = False):
    """latent [N,K] -> packed uint8 [N, ld_packed] (format C).  One kernel, one pass."""
    _cuda(w)
    assert w.dim() == 2
    w = w if w.stride(1) == 1 else w.contiguous()
    N, K = w.shape
    ld = packed_pitch(K) if ld_packed is None else int(ld_packed)
    packed = torch.empty((N, ld), dtype=torch.uint8, device=w.device)
    codes = torch.empty((N, K), dtype=torch.int8, device=w.device) if return_codes else None
    with torch.cuda.device(w.device):
        L.check(L.lib().tb200_quantize_pack(w.data_ptr(), _dt(w), N, K, w.stride(0), _MODES[mode],
                                            packed.data_ptr(), ld, _ptr(codes), _stream(w)), "quantize_pack")
    return (packed, codes) if return_codes else packed


def unpack(packed: torch.Tensor, K: int) -> torch.Tensor:
    """ops.unpack_ternary(pw, orig_dim): packed [N, ld] -> int8 [N, K]."""
    _cuda(packed)
    assert packed.dtype == torch.uint8 and packed.dim() == 2 and packed.stride(1) == 1
    N = packed.shape[0]
    out = torch.empty((N, K), dtype=torch.int8, device=packed.device)
    with torch.cuda.device(packed.device):
        L.check(L.lib().tb200_unpack(packed.data_ptr(), packed.stride(0), N, K, out.data_ptr(), _stream(packed)), "unpack")
    return out


def pack_format(codes: torch.Tensor, fmt: str) -> torch.Tensor:
    """int8 values in {-1,0,1} -> the reference's byte layout A, B (flattened, planar) or C ([N,K])."""
    _cuda(codes)
    codes = codes.to(torch.int8).contiguous()
    if fmt == "C":
        assert codes.dim() == 2
        N, K = codes.shape
        out = torch.empty((N, (K + 3) // 4), dtype=torch.uint8, device=codes.device)
    else:
        N, K = 1, codes.numel()
        out = torch.empty(((K + 3) // 4,), dtype=torch.uint8, device=codes.device)
    with torch.cuda.device(codes.device):
        L.check(L.lib().tb200_pack_format(codes.data_ptr(), N, K, _FMTS[fmt], out.data_ptr(), _stream(codes)), "pack_format")
    return out


def unpack_format(packed: torch.Tensor, fmt: str, N: int = 1, K: int | None = None) -> torch.Tensor:
    """Inverse of :func:`pack_format`.  A/B return the padded length 4*len(packed) like the reference."""
    _cuda(packed)
    packed = packed.contiguous()
    if fmt == "C":
        assert K is not None
        out = torch.empty((N, K), dtype=torch.int8, device=packed.device)
    else:
        N, K = 1, 4 * packed.numel()
        out = torch.empty((K,), dtype=torch.int8, device=packed.device)
    with torch.cuda.device(packed.device):
        L.check(L.lib().tb200_unpack_format(packed.data_ptr(), N, K, _FMTS[fmt], out.data_ptr(), _stream(packed)), "unpack_format")
    return out


def clip_latent_(w: torch.Tensor) -> torch.Tensor:
    """In-place clip(w, -1, 1): clip_binary_weights per tensor (binary/utils.py:74-77)."""
    _cuda(w)
    assert w.is_contiguous()
    with torch.cuda.device(w.device):
        L.check(L.lib().tb200_clip_latent(w.data_ptr(), _dt(w), w.numel(), _stream(w)), "clip_latent")
    return w


_train_ws: dict = {}


def _train_workspace(device) -> torch.Tensor:
    ws = _train_ws.get(device.index)
    if ws is None:
        ws = _train_ws[device.index] = torch.zeros(L.lib().tb200_train_workspace(), dtype=torch.uint8, device=device)
    return ws


def sumsq_accumulate_(acc: torch.Tensor, x: torch.Tensor, count_nonzero: bool = False) -> torch.Tensor:
    """``acc (0-dim float64, CUDA) += sum(x**2)`` or ``+= count_nonzero(x)`` (= ``sum(sign(x)**2)``), deterministic
    (``tb200_sumsq_accumulate``); no host synchronisation."""
    _cuda(acc, x)
    assert acc.dtype == torch.float64 and acc.numel() == 1 and x.is_contiguous()
    with torch.cuda.device(x.device):
        L.check(L.lib().tb200_sumsq_accumulate(x.data_ptr(), _dt(x), x.numel(), int(count_nonzero), acc.data_ptr(),
                                               _train_workspace(x.device).data_ptr(), _stream(x)), "sumsq_accumulate")
    return acc


def biobjective_apply_(grad: torch.Tensor, weights, norm_g_sq: torch.Tensor, norm_bg_sq: torch.Tensor) -> torch.Tensor:
    """In-place blend / rescale of one gradient tensor (``tb200_biobjective_apply``, binary/modules.py:166-176)."""
    _cuda(grad, weights, norm_g_sq, norm_bg_sq)
    assert grad.is_contiguous() and (weights is None or (weights.is_contiguous() and weights.dtype == grad.dtype and weights.numel() == grad.numel()))
    with torch.cuda.device(grad.device):
        L.check(L.lib().tb200_biobjective_apply(grad.data_ptr(), _ptr(weights), _dt(grad), grad.numel(), norm_g_sq.data_ptr(),
                                                norm_bg_sq.data_ptr(), _stream(grad)), "biobjective_apply")
    return grad


def ternary_init(weight: torch.Tensor, per_row: bool, iters: int = 10):
    """Ternary branch of linear_to_quantized (binary/modules.py:275-298) -> (latent f32, scale f32)."""
    _cuda(weight)
    w = weight.detach().contiguous()
    N, K = w.shape
    latent = torch.empty((N, K), dtype=torch.float32, device=w.device)
    scale = torch.empty((N,), dtype=torch.float32, device=w.device)
    ws = torch.empty((L.lib().tb200_ternary_init_workspace(N),), dtype=torch.uint8, device=w.device)
    with torch.cuda.device(w.device):
        L.check(L.lib().tb200_ternary_init(w.data_ptr(), _dt(w), N, K, int(per_row), int(iters),
                                           latent.data_ptr(), scale.data_ptr(), ws.data_ptr(), _stream(w)), "ternary_init")
    return latent, scale


def binary_init(weight: torch.Tensor, per_row: bool, positive_only: bool = False,
                prescale: float = 1.0, natural: bool = False):
    """Binary / binary01 branches of linear_to_quantized (binary/modules.py:193-266)."""
    _cuda(weight)
    w = weight.detach().contiguous()
    N, K = w.shape
    latent = torch.empty((N, K), dtype=torch.float32, device=w.device)
    scale = torch.empty((N,), dtype=torch.float32, device=w.device)
    ws = torch.empty((L.lib().tb200_ternary_init_workspace(N),), dtype=torch.uint8, device=w.device)
    with torch.cuda.device(w.device):
        L.check(L.lib().tb200_binary_init(w.data_ptr(), _dt(w), N, K, int(per_row), int(positive_only),
                                          float(prescale), int(natural), latent.data_ptr(), scale.data_ptr(),
                                          ws.data_ptr(), _stream(w)), "binary_init")
    return latent, scale


# ------------------------------------------------------------------ activations
def quant_act_i8(x: torch.Tensor, kperm: bool = False, layout: int | None = None):
    """x [M,K] -> (q_x int8 [M, ldq], s_x f32 [M], sum_qx i32 [M]).  layout 0 natural, 1 tcgen05
    GEMM operand order (``kperm=True``), 2 decode-GEMV image, 3 the field-split order of the 32-token list kernel;
    ldq = K rounded up to 16 (4096 for layout 2, 256 for layout 3)."""
    _cuda(x)
    assert x.dim() == 2 and x.stride(1) == 1
    M, K = x.shape
    layout = int(kperm) if layout is None else layout
    ldq = (K + 4095) // 4096 * 4096 if layout == 2 else (K + 255) // 256 * 256 if layout == 3 else (K + 15) // 16 * 16
    q = torch.empty((M, ldq), dtype=torch.int8, device=x.device)
    s = torch.empty((M,), dtype=torch.float32, device=x.device)
    sq = torch.empty((M,), dtype=torch.int32, device=x.device)
    with torch.cuda.device(x.device):
        L.check(L.lib().tb200_quant_act_i8(x.data_ptr(), _dt(x), M, K, x.stride(0), q.data_ptr(), ldq,
                                           s.data_ptr(), sq.data_ptr(), layout, _stream(x)), "quant_act_i8")
    return q, s, sq


# ------------------------------------------------------------------ the fused op
def _f32_or_none(t, n, device):
    if t is None:
        return None
    t = t.detach()
    if t.dtype != torch.float32 or not t.is_contiguous():
        t = t.to(torch.float32).contiguous()
    assert t.numel() == n and t.device == device
    return t


def gemv(x, packed, scale, bias, clip_val, out_dtype=None, debug=False, pdl=False, out=None):
    """Decode GEMV (M <= 32).  With ``debug`` also returns (acc i32 [M,N], s_x [M], q_x int8 [M,K])."""
    _cuda(x, packed, scale, bias)
    assert x.dim() == 2 and x.stride(1) == 1 and packed.dtype == torch.uint8
    M, K = x.shape
    N = packed.shape[0]
    out_dtype = out_dtype or x.dtype
    scale = _f32_or_none(scale, N, x.device)
    bias = _f32_or_none(bias, N, x.device)
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=x.device)
    acc = sx = qx = None
    if debug:
        acc = torch.empty((M, N), dtype=torch.int32, device=x.device)
        sx = torch.empty((M,), dtype=torch.float32, device=x.device)
        qx = torch.empty((M, K), dtype=torch.int8, device=x.device)
    with torch.cuda.device(x.device):
        L.check(L.lib().tb200_gemv(x.data_ptr(), _dt(x), M, K, x.stride(0), packed.data_ptr(), packed.stride(0), N,
                                   _ptr(scale), _ptr(bias), float(clip_val), out.data_ptr(), _DT[out.dtype],
                                   out.stride(0), _ptr(acc), _ptr(sx), _ptr(qx), L.FLAG_PDL if pdl else 0,
                                   _stream(x)), "gemv")
    return (out, acc, sx, qx) if debug else out


def gemv_q(q_x, s_x, sum_qx, K, packed, scale, bias, clip_val, out_dtype=torch.bfloat16, debug=False, pdl=False):
    """Decode GEMV on activations pre-quantised with ``quant_act_i8(layout=1)`` (the kperm order; 8 tokens per pass
    over the weights)."""
    _cuda(q_x, s_x, sum_qx, packed, scale, bias)
    M, N = q_x.shape[0], packed.shape[0]
    scale = _f32_or_none(scale, N, q_x.device)
    bias = _f32_or_none(bias, N, q_x.device)
    out = torch.empty((M, N), dtype=out_dtype, device=q_x.device)
    acc = torch.empty((M, N), dtype=torch.int32, device=q_x.device) if debug else None
    with torch.cuda.device(q_x.device):
        L.check(L.lib().tb200_gemv_q(q_x.data_ptr(), q_x.stride(0), s_x.data_ptr(), sum_qx.data_ptr(), M, K,
                                     packed.data_ptr(), packed.stride(0), N, _ptr(scale), _ptr(bias), float(clip_val),
                                     out.data_ptr(), _DT[out_dtype], out.stride(0), _ptr(acc),
                                     L.FLAG_PDL if pdl else 0, _stream(q_x)), "gemv_q")
    return (out, acc) if debug else out


def gemv_f32(x, packed, scale, bias, clip_val):
    """Float-activation GEMV (no activation quantisation): the reference's exact semantics."""
    _cuda(x, packed, scale, bias)
    assert x.dim() == 2 and x.stride(1) == 1
    M, K = x.shape
    N = packed.shape[0]
    scale = _f32_or_none(scale, N, x.device)
    bias = _f32_or_none(bias, N, x.device)
    out = torch.empty((M, N), dtype=x.dtype, device=x.device)
    with torch.cuda.device(x.device):
        L.check(L.lib().tb200_gemv_f32(x.data_ptr(), _dt(x), M, K, x.stride(0), packed.data_ptr(), packed.stride(0), N,
                                       _ptr(scale), _ptr(bias), float(clip_val), out.data_ptr(), _dt(out),
                                       out.stride(0), _stream(x)), "gemv_f32")
    return out


def gemm_i8(q_x, s_x, sum_qx, K, packed, scale, bias, clip_val, out_dtype=torch.bfloat16, debug=False, out=None):
    """tcgen05 GEMM on pre-quantised (kperm=1) activations.  ``debug`` also returns acc i32 [M,N]."""
    _cuda(q_x, s_x, sum_qx, packed, scale, bias)
    M = q_x.shape[0]
    N = packed.shape[0]
    scale = _f32_or_none(scale, N, q_x.device)
    bias = _f32_or_none(bias, N, q_x.device)
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=q_x.device)
    else:
        assert out.shape == (M, N) and out.stride(1) == 1
        out_dtype = out.dtype
    acc = torch.empty((M, N), dtype=torch.int32, device=q_x.device) if debug else None
    with torch.cuda.device(q_x.device):
        L.check(L.lib().tb200_gemm_i8(q_x.data_ptr(), q_x.stride(0), s_x.data_ptr(), sum_qx.data_ptr(), M, K,
                                      packed.data_ptr(), packed.stride(0), N, _ptr(scale), _ptr(bias),
                                      float(clip_val), out.data_ptr(), _DT[out_dtype], out.stride(0), _ptr(acc),
                                      _stream(q_x)), "gemm_i8")
    return (out, acc) if debug else out


_ws_cache: dict = {}
_ws_retired: list = []


def _workspace(M: int, K: int, device) -> torch.Tensor:
    """Per-(device, stream) scratch for the activation quantiser.  A workspace that has been handed out is NEVER freed:
    a CUDA graph captured through ternary_matmul has its address baked into its kernel nodes, so outgrown generations
    are retired to a keep-alive list instead of being released (growth is geometric, so at most ~2x the largest need
    stays allocated).  While a stream is capturing, a workspace private to that capture is allocated from the graph's
    own memory pool, so replays never alias an eager call's scratch."""
    need = L.lib().tb200_linear_workspace(M, K)
    stream = torch.cuda.current_stream(device)
    if torch.cuda.is_current_stream_capturing():
        return torch.empty((max(need, 256),), dtype=torch.uint8, device=device)      # owned by the graph's private pool
    key = (device.index, stream.cuda_stream)
    ws = _ws_cache.get(key)
    if ws is None or ws.numel() < need:
        if ws is not None:
            _ws_retired.append(ws)
        ws = torch.empty((max(need, 1 << 20, 2 * (ws.numel() if ws is not None else 0)),), dtype=torch.uint8, device=device)
        _ws_cache[key] = ws
    return ws


def ternary_matmul(x: torch.Tensor, pw: torch.Tensor, scale, bias, clip_val: float, out_dtype=None, out=None) -> torch.Tensor:
    """``ops.ternary_matmul(x, pw, scale, bias, clip_val)`` (impl/tf/moondream/ternary.py:279-281):
    ``clip((x @ unpack(pw)^T) * scale + bias, +-clip_val)`` for ``x [..., K]`` and ``pw [N, ld]``.
    M <= 8 tokens: one GEMV launch with the activation quantisation fused in its prologue; M <= 16: quantiser +
    two 8-token GEMV passes; larger M: quantiser + the tcgen05 kind::i8 GEMM (``tb200_linear_forward``).
    Under ``torch.compile`` the call is routed through the registered custom op ``ternary_b200::ternary_matmul``
    (an opaque node: no graph break)."""
    if out is None and torch.compiler.is_compiling():
        return torch.ops.ternary_b200.ternary_matmul(x, pw, None if scale is None else scale.detach(),
                                                     None if bias is None else bias.detach(), float(clip_val),
                                                     x.dtype if out_dtype is None else out_dtype)
    _cuda(x, pw, scale, bias)
    K = x.shape[-1]
    N = pw.shape[0]
    x2 = x.reshape(-1, K)
    if x2.stride(-1) != 1:
        x2 = x2.contiguous()
    M = x2.shape[0]
    out_dtype = out_dtype or x.dtype
    scale = _f32_or_none(scale, N, x.device)
    bias = _f32_or_none(bias, N, x.device)
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=x.device)
    else:
        assert out.shape == (M, N) and out.stride(1) == 1
        out_dtype = out.dtype
    if M > 0:
        ws = _workspace(M, K, x.device)
        with torch.cuda.device(x.device):
            L.check(L.lib().tb200_linear_forward(x2.data_ptr(), _dt(x2), M, K, x2.stride(0), pw.data_ptr(),
                                                 pw.stride(0), N, _ptr(scale), _ptr(bias), float(clip_val),
                                                 out.data_ptr(), _DT[out_dtype], out.stride(0), ws.data_ptr(),
                                                 _stream(x)), "linear_forward")
    return out.reshape(*x.shape[:-1], N)


def gemv_stream(xs, packed, scale, bias, clip_val, out=None, out_dtype=None, debug=False, workspace=None):
    """Streaming decode over a list of layers in one persistent launch (``tb200_gemv_stream``).
    ``xs [R, M, K]`` (or ``[M, K]`` shared by all layers), ``packed [R, N, ld]``, ``scale``/``bias`` ``[R, N]`` or None
    -> ``out [R, M, N]``.  ``debug`` also returns the integer accumulators of layer 0."""
    _cuda(xs, packed, scale, bias)
    assert packed.dim() == 3 and packed.dtype == torch.uint8 and packed.stride(2) == 1
    R, N, _ = packed.shape
    shared_x = xs.dim() == 2
    assert xs.stride(-1) == 1
    M, K = xs.shape[-2], xs.shape[-1]
    out_dtype = out_dtype or xs.dtype
    if out is None:
        out = torch.empty((R, M, N), dtype=out_dtype, device=xs.device)
    assert out.shape == (R, M, N) and out.stride(2) == 1
    if scale is not None:
        assert scale.dtype == torch.float32 and scale.shape == (R, N) and scale.stride(1) == 1
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.shape == (R, N) and bias.stride(1) == 1
    stride_sb = scale.stride(0) if scale is not None else (bias.stride(0) if bias is not None else 0)
    if scale is not None and bias is not None:
        assert scale.stride(0) == bias.stride(0)
    need = L.lib().tb200_gemv_stream_workspace(M, K, R)
    if workspace is None:
        workspace = torch.empty((need,), dtype=torch.uint8, device=xs.device)
    assert workspace.numel() >= need
    acc = torch.empty((M, N), dtype=torch.int32, device=xs.device) if debug else None
    with torch.cuda.device(xs.device):
        L.check(L.lib().tb200_gemv_stream(xs.data_ptr(), _dt(xs), M, K, xs.stride(-2), 0 if shared_x else xs.stride(0),
                                          packed.data_ptr(), packed.stride(1), packed.stride(0), N,
                                          _ptr(scale), _ptr(bias), stride_sb, float(clip_val),
                                          out.data_ptr(), _DT[out.dtype], out.stride(1), out.stride(0), R,
                                          _ptr(acc), workspace.data_ptr(), _stream(xs)), "gemv_stream")
    return (out, acc) if debug else out


def gemv_stream_tp(xs, packed, scale, bias, clip_val, peer_ptrs, n_peers: int, out_dtype, ldo: int, stride_out: int,
                   workspace=None):
    """Tensor-parallel ``gemv_stream`` (``tb200_gemv_stream_tp``): ``packed [R, N_local, ld]`` is this rank's row block
    of every layer; the epilogues store the rank's ``[R, M, N_local]`` outputs into every rank's ``[R, M, ldo]`` buffer.
    ``peer_ptrs``: ctypes array of ``n_peers`` addresses, each already offset to this rank's first column.  The
    caller issues the cross-rank barrier."""
    _cuda(xs, packed, scale, bias)
    assert packed.dim() == 3 and packed.dtype == torch.uint8 and packed.stride(2) == 1
    R, N, _ = packed.shape
    shared_x = xs.dim() == 2
    assert xs.stride(-1) == 1
    M, K = xs.shape[-2], xs.shape[-1]
    if scale is not None:
        assert scale.dtype == torch.float32 and scale.shape == (R, N) and scale.stride(1) == 1
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.shape == (R, N) and bias.stride(1) == 1
    stride_sb = scale.stride(0) if scale is not None else (bias.stride(0) if bias is not None else 0)
    if scale is not None and bias is not None:
        assert scale.stride(0) == bias.stride(0)
    need = L.lib().tb200_gemv_stream_workspace(M, K, R)
    if workspace is None:
        workspace = torch.empty((need,), dtype=torch.uint8, device=xs.device)
    assert workspace.numel() >= need
    with torch.cuda.device(xs.device):
        L.check(L.lib().tb200_gemv_stream_tp(xs.data_ptr(), _dt(xs), M, K, xs.stride(-2), 0 if shared_x else xs.stride(0),
                                             packed.data_ptr(), packed.stride(1), packed.stride(0), N,
                                             _ptr(scale), _ptr(bias), stride_sb, float(clip_val),
                                             peer_ptrs, n_peers, _DT[out_dtype], ldo, stride_out, R,
                                             workspace.data_ptr(), _stream(xs)), "gemv_stream_tp")


def gemv_stream_mc(xs, packed, scale, bias, clip_val, out_multicast: int, out_dtype, ldo: int, stride_out: int, workspace=None):
    """``gemv_stream_tp`` through an NVSwitch multicast mapping (``tb200_gemv_stream_mc``): ``out_multicast`` is the multicast
    address of this rank's first column in the symmetric ``[R, M, ldo]`` buffers."""
    _cuda(xs, packed, scale, bias)
    assert packed.dim() == 3 and packed.dtype == torch.uint8 and packed.stride(2) == 1
    R, N, _ = packed.shape
    shared_x = xs.dim() == 2
    assert xs.stride(-1) == 1
    M, K = xs.shape[-2], xs.shape[-1]
    stride_sb = scale.stride(0) if scale is not None else (bias.stride(0) if bias is not None else 0)
    need = L.lib().tb200_gemv_stream_workspace(M, K, R)
    if workspace is None:
        workspace = torch.empty((need,), dtype=torch.uint8, device=xs.device)
    assert workspace.numel() >= need
    with torch.cuda.device(xs.device):
        L.check(L.lib().tb200_gemv_stream_mc(xs.data_ptr(), _dt(xs), M, K, xs.stride(-2), 0 if shared_x else xs.stride(0),
                                             packed.data_ptr(), packed.stride(1), packed.stride(0), N,
                                             _ptr(scale), _ptr(bias), stride_sb, float(clip_val),
                                             int(out_multicast), _DT[out_dtype], ldo, stride_out, R,
                                             workspace.data_ptr(), _stream(xs)), "gemv_stream_mc")


def launch_count() -> int:
    return int(L.lib().tb200_launch_count())


# ------------------------------------------------------------------ 1-bit packed weights (binary / binary01 layers)
def b1_pitch(K: int) -> int:
    return int(L.lib().tb200_b1_pitch(K))


def quantize_pack_b1(w: torch.Tensor) -> torch.Tensor:
    """Latent weights ``[N, K]`` -> format B1 ``[N, 16 * ceil(K/128)]`` uint8: bit ``k & 7`` of byte ``k >> 3`` is
    ``w[n, k] >= 0`` (sign_ste / step_ste, binary/modules.py:5-18)."""
    _cuda(w)
    assert w.dim() == 2
    w = w.detach()
    if w.stride(1) != 1:
        w = w.contiguous()
    N, K = w.shape
    out = torch.empty((N, b1_pitch(K)), dtype=torch.uint8, device=w.device)
    with torch.cuda.device(w.device):
        L.check(L.lib().tb200_quantize_pack_b1(w.data_ptr(), _dt(w), N, K, w.stride(0), out.data_ptr(), out.stride(0), _stream(w)),
                "quantize_pack_b1")
    return out


def expand_b1(packed1: torch.Tensor, K: int, binary01: bool) -> torch.Tensor:
    """B1 -> the 2-bit format C ``[N, packed_pitch(K)]`` (codes value + 1) the 2-bit kernels read."""
    _cuda(packed1)
    N = packed1.shape[0]
    out = torch.empty((N, packed_pitch(K)), dtype=torch.uint8, device=packed1.device)
    with torch.cuda.device(packed1.device):
        L.check(L.lib().tb200_expand_b1(packed1.data_ptr(), packed1.stride(0), N, K, int(binary01), out.data_ptr(), out.stride(0),
                                        _stream(packed1)), "expand_b1")
    return out


B1_GEMV_MAX_M = 2


def binary_matmul(x: torch.Tensor, pw1: torch.Tensor, K: int, scale, bias, clip_val: float, binary01: bool, out_dtype=None,
                  force_b1: bool = False, pdl: bool = False):
    """The layer forward on 1-bit weights: ``clip((x @ W^T) * scale + bias)`` with ``W = 2 bit - 1`` (binary) or ``bit``
    (binary01).  Up to 2 tokens: ONE launch of ``tb200_gemv_b1_fused`` (half the weight bytes of the 2-bit GEMV); more: the
    weights are expanded to format C on the fly and go through ``tb200_linear_forward``.  ``force_b1`` (tests): quantiser +
    ``tb200_gemv_b1`` for up to 8 tokens."""
    _cuda(x, pw1, scale, bias)
    assert x.shape[-1] == K and pw1.dtype == torch.uint8 and pw1.dim() == 2 and pw1.stride(1) == 1
    N = pw1.shape[0]
    x2 = x.reshape(-1, K)
    if x2.stride(-1) != 1:
        x2 = x2.contiguous()
    M = x2.shape[0]
    out_dtype = out_dtype or x.dtype
    scale = _f32_or_none(scale, N, x.device)
    bias = _f32_or_none(bias, N, x.device)
    if M > B1_GEMV_MAX_M and not force_b1:
        # dp4a computes one token per instruction (mma.sync eight): above two tokens the 2-bit kernels win even after the
        # expansion pass
        y = ternary_matmul(x2, expand_b1(pw1, K, binary01), scale, bias, clip_val, out_dtype=out_dtype)
        return y.reshape(*x.shape[:-1], N)
    out = torch.empty((M, N), dtype=out_dtype, device=x.device)
    if 0 < M <= 2 and not force_b1:
        with torch.cuda.device(x.device):
            L.check(L.lib().tb200_gemv_b1_fused(x2.data_ptr(), _dt(x2), M, K, x2.stride(0), pw1.data_ptr(), pw1.stride(0), N,
                                                _ptr(scale), _ptr(bias), float(clip_val), int(binary01), out.data_ptr(),
                                                _DT[out_dtype], out.stride(0), L.FLAG_PDL if pdl else 0, _stream(x)), "gemv_b1_fused")
    elif M > 0:
        q, s, sq = quant_act_i8(x2, layout=0)
        with torch.cuda.device(x.device):
            L.check(L.lib().tb200_gemv_b1(q.data_ptr(), q.stride(0), s.data_ptr(), sq.data_ptr(), M, K, pw1.data_ptr(), pw1.stride(0), N,
                                          _ptr(scale), _ptr(bias), float(clip_val), int(binary01), out.data_ptr(), _DT[out_dtype],
                                          out.stride(0), L.FLAG_PDL, _stream(x)), "gemv_b1")
    return out.reshape(*x.shape[:-1], N)


# ------------------------------------------------------------------ torch.library registration
# The C-ABI calls above are invisible to torch.compile / export; the ops below wrap them as opaque custom ops with
# shape-only ("fake") implementations, so a compiled model keeps the layer as ONE node instead of graph-breaking.
@torch.library.custom_op("ternary_b200::ternary_matmul", mutates_args=(), device_types="cuda")
def _ternary_matmul_op(x: torch.Tensor, pw: torch.Tensor, scale: torch.Tensor | None, bias: torch.Tensor | None,
                       clip_val: float, out_dtype: torch.dtype) -> torch.Tensor:
    return ternary_matmul(x, pw, scale, bias, clip_val, out_dtype=out_dtype)


@_ternary_matmul_op.register_fake
def _(x, pw, scale, bias, clip_val, out_dtype):
    return x.new_empty((*x.shape[:-1], pw.shape[0]), dtype=out_dtype)


@torch.library.custom_op("ternary_b200::quantize_pack", mutates_args=(), device_types="cuda")
def _quantize_pack_op(w: torch.Tensor, mode: str) -> torch.Tensor:
    return quantize_pack(w, mode)


@_quantize_pack_op.register_fake
def _(w, mode):
    return w.new_empty((w.shape[0], packed_pitch(w.shape[1])), dtype=torch.uint8)


@torch.library.custom_op("ternary_b200::unpack", mutates_args=(), device_types="cuda")
def _unpack_op(packed: torch.Tensor, K: int) -> torch.Tensor:
    return unpack(packed, K)


@_unpack_op.register_fake
def _(packed, K):
    return packed.new_empty((packed.shape[0], K), dtype=torch.int8)


@torch.library.custom_op("ternary_b200::quant_act_i8", mutates_args=(), device_types="cuda")
def _quant_act_op(x: torch.Tensor, layout: int) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    return quant_act_i8(x, layout=layout)


@_quant_act_op.register_fake
def _(x, layout):
    M, K = x.shape
    ldq = (K + 4095) // 4096 * 4096 if layout == 2 else (K + 255) // 256 * 256 if layout == 3 else (K + 15) // 16 * 16
    return (x.new_empty((M, ldq), dtype=torch.int8), x.new_empty((M,), dtype=torch.float32),
            x.new_empty((M,), dtype=torch.int32))


# ------------------------------------------------------------------ a chain of dependent layers (fused MLP block)
_ACTS = {None: L.ACT_NONE, "none": L.ACT_NONE, "gelu_new": L.ACT_GELU_TANH, "gelu_tanh": L.ACT_GELU_TANH,
         "gelu_pytorch_tanh": L.ACT_GELU_TANH, "relu": L.ACT_RELU}


class ChainPlan:
    """Plan + workspace of ``tb200_chain_forward``: ``layers`` is a list of ``(packed [N, ld] uint8, scale [N] | None,
    bias [N] | None, clip_val, activation)`` with ``activation`` in {None, "gelu_new", "relu"}; every layer consumes the
    previous one's output (``K_l == N_{l-1}``).  The reference wires fc1 -> gelu_new -> fc2 this way
    (impl/tf/moondream/layers.py:32-43; the activation is applied after the clip, ternary.py:245-246).
    The plan keeps references to the tensors it points to.  Decode shapes only: ``M <= 8`` tokens per call."""

    def __init__(self, layers, in_features: int, dtype=torch.bfloat16, max_tokens: int = L.CHAIN_MAX_M):
        assert 1 <= len(layers) <= L.CHAIN_MAX_LAYERS and 1 <= max_tokens <= L.CHAIN_MAX_M
        self.dtype, self.max_tokens, self.in_features = dtype, max_tokens, in_features
        dev = layers[0][0].device
        self._keep = []
        descs = (L.ChainLayerDesc * len(layers))()
        K = in_features
        for i, (pw, scale, bias, clip_val, act) in enumerate(layers):
            _cuda(pw, scale, bias)
            assert pw.dtype == torch.uint8 and pw.dim() == 2 and pw.stride(1) == 1
            N = pw.shape[0]
            scale, bias = _f32_or_none(scale, N, dev), _f32_or_none(bias, N, dev)
            self._keep += [pw, scale, bias]
            descs[i] = L.ChainLayerDesc(pw.data_ptr(), pw.stride(0), N, K, _ptr(scale), _ptr(bias), float(clip_val), _ACTS[act])
            K = N
        self.out_features = K
        nbytes = L.lib().tb200_chain_plan_bytes(len(layers))
        host = torch.zeros(nbytes, dtype=torch.uint8)
        ws_bytes = L.C.c_size_t(0)
        esz = torch.empty((), dtype=dtype).element_size()
        with torch.cuda.device(dev):
            L.check(L.lib().tb200_chain_plan_build(descs, len(layers), max_tokens, esz, host.data_ptr(), L.C.byref(ws_bytes)),
                    "chain_plan_build")
        self.plan_host = host
        self.plan_dev = host.to(dev)                          # cudaMalloc'ed: 256-byte aligned
        self.workspace = torch.zeros(ws_bytes.value, dtype=torch.uint8, device=dev)     # control block zero before first use
        self.n_layers = len(layers)

    def hidden(self, layer: int, M: int) -> torch.Tensor:
        """The inter-layer activations the LAST call left in the workspace: output of ``layer`` (not the last one),
        ``[M, N_layer]`` in the plan's dtype.  The kernel stores them bit-inverted (an all-zero element means "not written")
        in the buffer set of the call's generation parity.  For parity tests."""
        import struct
        base = 64 + 192 * layer
        raw = bytes(self.plan_host[base + 128: base + 192].tolist())
        N, = struct.unpack_from("<i", raw, 16)
        pitch, = struct.unpack_from("<i", raw, 44)
        off, set_stride = struct.unpack_from("<qq", raw, 48)
        gen = int(self.workspace[:4].view(torch.int32).item())            # generation AFTER the call
        esz = torch.empty((), dtype=self.dtype).element_size()
        start = off + ((gen - 1) & 1) * set_stride
        ity = torch.int32 if esz == 4 else torch.int16
        enc = self.workspace[start: start + M * pitch * esz].view(ity).view(M, pitch)[:, :N]
        assert bool((enc != 0).all()), "element not written"
        return (~enc).view(self.dtype)

    def forward(self, x: torch.Tensor, out: torch.Tensor | None = None, out_dtype=None, pdl: bool = False) -> torch.Tensor:
        _cuda(x)
        assert x.dtype == self.dtype and x.shape[-1] == self.in_features
        x2 = x.reshape(-1, self.in_features)
        if x2.stride(-1) != 1:
            x2 = x2.contiguous()
        M = x2.shape[0]
        if M > self.max_tokens:
            raise TernaryB200Error(f"ChainPlan serves at most {self.max_tokens} tokens per call, got {M}")
        if out is None:
            out = torch.empty((M, self.out_features), dtype=out_dtype or x.dtype, device=x.device)
        with torch.cuda.device(x.device):
            L.check(L.lib().tb200_chain_forward(self.plan_host.data_ptr(), self.plan_dev.data_ptr(), x2.data_ptr(), _dt(x2), M,
                                                x2.stride(0), out.data_ptr(), _DT[out.dtype], out.stride(0),
                                                self.workspace.data_ptr(), L.FLAG_PDL if pdl else 0, _stream(x)), "chain_forward")
        return out.reshape(*x.shape[:-1], self.out_features)
