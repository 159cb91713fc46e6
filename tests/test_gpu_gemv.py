"""GPU parity, decode GEMV: integer accumulators bit-exact vs the oracle; float outputs within
1e-3 of the reference layer fed the de-quantised activations (SURVEY.md §7 parity definition)."""
import numpy as np
import pytest
import torch

from oracle import ref_int8, ref_modules as om
from ternary_vllm_b200 import ops

pytestmark = pytest.mark.gpu

TOL = 1e-3     # BASELINE.json: max rel err 1e-3 (relative to max(|ref|, rms(ref)))


def make(N, K, M, seed, xdtype=torch.bfloat16, scale_mode="sqrt"):
    g = torch.Generator().manual_seed(seed)
    W = torch.randn(N, K, generator=g).clip(-1, 1)
    x = torch.randn(M, K, generator=g).to(xdtype)
    bias = torch.randn(N, generator=g)
    scale = torch.full((N,), 1.0 / K ** 0.5) if scale_mode == "sqrt" else torch.ones(N)
    return W, x, scale, bias


def check_against_oracle(W, x, scale, bias, clip, out, acc, sx, qx, mode="ternary"):
    codes = ref_int8.codes_of(W.numpy(), mode)
    qo, so, _ = ref_int8.quant_act(x.float().numpy())
    assert np.array_equal(qx.cpu().numpy(), qo), "activation codes differ"
    assert np.array_equal(sx.cpu().numpy(), so), "activation scales differ"
    acc_o = ref_int8.int_acc(qo, codes)
    assert np.array_equal(acc.cpu().numpy().astype(np.int64), acc_o), "integer accumulators differ"
    y_o = ref_int8.fused_out(acc_o, so, None if scale is None else scale.numpy(), None if bias is None else bias.numpy(), clip)
    got = out.float().cpu().numpy()
    if out.dtype == torch.float32:
        assert np.allclose(got, y_o, rtol=2e-6, atol=1e-6)
    else:   # one rounding to the 16-bit output type
        want = torch.from_numpy(y_o).to(out.dtype).float().numpy()
        ulp = np.abs(want) * (2.0 ** -7 if out.dtype == torch.bfloat16 else 2.0 ** -10) + 1e-6
        assert np.all(np.abs(got - want) <= ulp)
    # tie to the reference: its fake-quant layer on the de-quantised activations
    xd = torch.from_numpy(qo.astype(np.float32) * so[:, None])
    y_ref = om.scaled_linear_forward(xd, W, scale if scale is not None else torch.ones(W.shape[0]), bias, clip, mode).numpy()
    tol = TOL * np.maximum(np.abs(y_ref), np.sqrt((y_ref ** 2).mean()))
    y32 = ref_int8.fused_out(acc.cpu().numpy(), sx.cpu().numpy(), None if scale is None else scale.numpy(),
                             None if bias is None else bias.numpy(), clip)
    assert np.all(np.abs(y32 - y_ref) <= tol)


@pytest.mark.parametrize("M", [1, 2, 3, 4, 5, 8, 9, 17, 32])
@pytest.mark.parametrize("N,K", [(64, 64), (48, 100), (37, 259), (512, 2048), (1000, 4096), (130, 11008)])
def test_gemv_parity(cuda, M, N, K):
    W, x, scale, bias = make(N, K, M, seed=M * 100003 + N * 17 + K)
    packed = ops.quantize_pack(W.to(cuda), "ternary")
    out, acc, sx, qx = ops.gemv(x.to(cuda), packed, scale.to(cuda), bias.to(cuda), 100.0, debug=True)
    assert out.dtype == torch.bfloat16 and out.shape == (M, N)
    check_against_oracle(W, x, scale, bias, 100.0, out, acc, sx, qx)


@pytest.mark.parametrize("M", [3, 4, 7, 8, 9, 16, 17, 32])
@pytest.mark.parametrize("N,K", [(48, 100), (512, 2048), (1000, 4096), (130, 11008), (2000, 5000), (40, 20000)])
def test_gemv_prequantised_parity(cuda, M, N, K):
    """tb200_quant_act_i8(layout 1) + tb200_gemv_q: the path tb200_linear_forward takes for 3..32 tokens."""
    W, x, scale, bias = make(N, K, M, seed=M * 31 + N + K)
    packed = ops.quantize_pack(W.to(cuda), "ternary")
    q, s, sq = ops.quant_act_i8(x.to(cuda), layout=1)
    out, acc = ops.gemv_q(q, s, sq, K, packed, scale.to(cuda), bias.to(cuda), 100.0, debug=True, pdl=True)
    qo, so, sqo = ref_int8.quant_act(x.float().numpy())
    assert np.array_equal(s.cpu().numpy(), so) and np.array_equal(sq.cpu().numpy(), sqo)
    acc_o = ref_int8.int_acc(qo, ref_int8.codes_of(W.numpy()))
    assert np.array_equal(acc.cpu().numpy().astype(np.int64), acc_o)
    y_o = ref_int8.fused_out(acc_o, so, scale.numpy(), bias.numpy(), 100.0)
    want = torch.from_numpy(y_o).bfloat16().float().numpy()
    assert np.all(np.abs(out.float().cpu().numpy() - want) <= np.abs(want) * 2.0 ** -7 + 1e-6)
    # and the public op (quantise + GEMV chosen by token count) gives the same bits
    y = ops.ternary_matmul(x.to(cuda), packed, scale.to(cuda), bias.to(cuda), 100.0)
    assert torch.equal(y, out)


def test_gemv_grouped_matches_single(cuda):
    R, N, K = 5, 300, 1024
    g = torch.Generator().manual_seed(8)
    packs, outs = [], []
    x = torch.randn(R, 2, K, generator=g).bfloat16().to(cuda)
    scale = (torch.rand(R, N, generator=g) * 0.1).to(cuda)
    bias = torch.randn(R, N, generator=g).to(cuda)
    for r in range(R):
        packs.append(ops.quantize_pack(torch.randn(N, K, generator=g).clip(-1, 1).to(cuda), "ternary"))
        outs.append(ops.gemv(x[r], packs[r], scale[r], bias[r], 100.0))
    stacked = torch.stack(packs)
    from ternary_vllm_b200 import _lib as L
    out = torch.empty(R, 2, N, dtype=torch.bfloat16, device=cuda)
    L.check(L.lib().tb200_gemv_grouped(x.data_ptr(), L.TB200_BF16, 2, K, K, 2 * K, stacked.data_ptr(), stacked.stride(1),
                                       stacked.stride(0), N, scale.data_ptr(), bias.data_ptr(), N, 100.0, out.data_ptr(),
                                       L.TB200_BF16, N, 2 * N, R, 0, torch.cuda.current_stream().cuda_stream), "gemv_grouped")
    torch.cuda.synchronize()
    assert torch.equal(out, torch.stack(outs))


@pytest.mark.parametrize("xdtype,odtype", [(torch.float32, torch.float32), (torch.float16, torch.float16),
                                           (torch.bfloat16, torch.float32), (torch.float16, torch.float32)])
def test_gemv_dtypes_and_options(cuda, xdtype, odtype):
    W, x, scale, bias = make(200, 1024, 3, seed=9, xdtype=xdtype)
    packed = ops.quantize_pack(W.to(cuda), "ternary")
    out, acc, sx, qx = ops.gemv(x.to(cuda), packed, scale.to(cuda), bias.to(cuda), 100.0, out_dtype=odtype, debug=True)
    check_against_oracle(W, x, scale, bias, 100.0, out, acc, sx, qx)
    out, acc, sx, qx = ops.gemv(x.to(cuda), packed, None, None, 100.0, out_dtype=odtype, debug=True)     # no scale, no bias
    check_against_oracle(W, x, None, None, 100.0, out, acc, sx, qx)


def test_gemv_clip_is_live(cuda):
    W, x, scale, bias = make(512, 2048, 4, seed=3, xdtype=torch.float32, scale_mode="one")
    packed = ops.quantize_pack(W.to(cuda), "ternary")
    out, acc, sx, qx = ops.gemv(x.to(cuda), packed, scale.to(cuda), bias.to(cuda), 100.0, debug=True)
    check_against_oracle(W, x, scale, bias, 100.0, out, acc, sx, qx)
    assert (out.abs() == 100.0).float().mean() > 0.001          # ~0.6% of outputs clip at K=2048


@pytest.mark.parametrize("mode", ["binary", "binary01"])
def test_gemv_binary_modes(cuda, mode):
    W, x, scale, bias = make(96, 320, 2, seed=4, xdtype=torch.float32)
    packed = ops.quantize_pack(W.to(cuda), mode)
    out, acc, sx, qx = ops.gemv(x.to(cuda), packed, scale.to(cuda), bias.to(cuda), 100.0, debug=True)
    check_against_oracle(W, x, scale, bias, 100.0, out, acc, sx, qx, mode)


def test_gemv_zero_and_empty(cuda):
    W, x, scale, bias = make(32, 64, 2, seed=5, xdtype=torch.float32)
    x[0] = 0
    packed = ops.quantize_pack(W.to(cuda), "ternary")
    out, acc, sx, qx = ops.gemv(x.to(cuda), packed, scale.to(cuda), bias.to(cuda), 100.0, debug=True)
    check_against_oracle(W, x, scale, bias, 100.0, out, acc, sx, qx)
    assert torch.all(acc[0] == 0)
    e = ops.gemv(torch.empty(0, 64, device=cuda), packed, None, None, 100.0)
    assert e.shape == (0, 32)


def test_gemv_pdl_chain_matches_plain(cuda):
    """Back-to-back layers under programmatic dependent launch give the same bits."""
    W1, x, s1, b1 = make(1024, 1024, 1, seed=6)
    W2, _, s2, b2 = make(1024, 1024, 1, seed=7)
    p1, p2 = ops.quantize_pack(W1.to(cuda), "ternary"), ops.quantize_pack(W2.to(cuda), "ternary")
    xg = x.to(cuda)
    torch.cuda.synchronize()
    y = xg
    for _ in range(8):
        y = ops.gemv(y, p1, s1.to(cuda), b1.to(cuda), 100.0)
        y = ops.gemv(y, p2, s2.to(cuda), b2.to(cuda), 100.0)
    z = xg
    for _ in range(8):
        z = ops.gemv(z, p1, s1.to(cuda), b1.to(cuda), 100.0, pdl=True)
        z = ops.gemv(z, p2, s2.to(cuda), b2.to(cuda), 100.0, pdl=True)
    torch.cuda.synchronize()
    assert torch.equal(y, z)


@pytest.mark.parametrize("tag", ["small", "k4", "ragged"])
def test_gemv_f32_matches_reference_golden(cuda, golden, tag):
    """Float-activation GEMV vs the live reference's recorded outputs (no activation quantisation)."""
    W, x = torch.from_numpy(golden[f"{tag}_W"]), torch.from_numpy(golden[f"{tag}_x"])
    scale, bias = torch.from_numpy(golden[f"{tag}_scale"]), torch.from_numpy(golden[f"{tag}_bias"])
    for mode, key in (("ternary", "y_ternary"), ("binary", "y_binary"), ("binary01", "y_binary01")):
        packed = ops.quantize_pack(W.to(cuda), mode)
        y = ops.gemv_f32(x.to(cuda), packed, scale.to(cuda), bias.to(cuda), 100.0).cpu().numpy()
        ref = golden[f"{tag}_{key}"]
        assert np.all(np.abs(y - ref) <= 1e-4 * np.maximum(np.abs(ref), np.sqrt((ref ** 2).mean())))
    packed = ops.quantize_pack(W.to(cuda), "ternary")
    y = ops.gemv_f32(x.to(cuda), packed, (torch.ones_like(scale) * 4).to(cuda), bias.to(cuda), 10.0).cpu().numpy()
    assert np.allclose(y, golden[f"{tag}_y_ternary_clip"], rtol=1e-4, atol=1e-4)
    y = ops.gemv_f32(x.to(cuda), packed, scale.to(cuda), None, 100.0).cpu().numpy()
    assert np.allclose(y, golden[f"{tag}_y_ternary_nobias"], rtol=1e-4, atol=1e-4)


def test_config0_2048_vs_reference_golden(cuda, golden):
    """BASELINE.json configs[0]: 2048x2048 ternary Linear, batch 1, fp32, vs the live reference's output."""
    assert str(golden["torch_version"]) == torch.__version__
    torch.manual_seed(1234)
    W = torch.randn(2048, 2048).clip(-1, 1)
    x = torch.randn(1, 2048)
    bias = torch.randn(2048)
    scale = torch.full((2048,), 1.0 / 2048 ** 0.5)
    packed, codes = ops.quantize_pack(W.to(cuda), "ternary", return_codes=True)
    assert [(codes == v).sum().item() for v in (-1, 0, 1)] == golden["cfg0_code_hist"].tolist()
    ref = golden["cfg0_y"]
    rms = np.sqrt((ref ** 2).mean())
    y = ops.gemv_f32(x.to(cuda), packed, scale.to(cuda), bias.to(cuda), 100.0).cpu().numpy()
    assert np.all(np.abs(y - ref) <= 1e-4 * np.maximum(np.abs(ref), rms))          # exact-semantics path
    # int8-activation path: bit-exact accumulators, and the end-to-end error vs raw activations is REPORTED
    out, acc, sx, qx = ops.gemv(x.to(cuda), packed, scale.to(cuda), bias.to(cuda), 100.0, debug=True)
    check_against_oracle(W, x, scale, bias, 100.0, out, acc, sx, qx)
    err = np.abs(out.cpu().numpy() - ref).max() / rms
    print(f"config0 int8-activation error vs float-activation reference: {err:.4f} of rms")
    assert err < 0.05
    y1 = ops.gemv_f32(x.to(cuda), packed, torch.ones(2048, device=cuda), bias.to(cuda), 100.0).cpu().numpy()
    assert np.allclose(y1, golden["cfg0_y_scale1"], rtol=1e-4, atol=1e-3)
