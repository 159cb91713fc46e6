"""GPU parity of the 1-bit packed path of the two-valued layers (ScaledBinaryLinear,
ScaledBinary01Linear, SmoothlyBinarizedLinear; binary/modules.py:55-90, 112-146): packing bit-exact against numpy,
integer arithmetic bit-exact against oracle/ref_int8, outputs against the oracle, the 2-bit path and the live reference's
recorded forwards."""
import numpy as np
import pytest
import torch

import ternary_vllm_b200 as tb
from oracle import ref_int8, ref_modules as om
from ternary_vllm_b200 import ops

pytestmark = pytest.mark.gpu
TOL = 1e-3


def _w(N, K, seed, dtype=torch.float32):
    g = torch.Generator().manual_seed(seed)
    W = torch.randn(N, K, generator=g)
    W[0, : min(K, 4)] = torch.tensor([0.0, -0.0, 1e-30, -1e-30])[: min(K, 4)]      # the decision boundary: -0.0 >= 0 is true
    return W.to(dtype)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("N,K", [(5, 8), (33, 100), (64, 128), (17, 300), (128, 4096), (9, 4500)])
def test_quantize_pack_b1_bit_exact(cuda, N, K, dtype):
    W = _w(N, K, N + K, dtype)
    got = ops.quantize_pack_b1(W.to(cuda)).cpu().numpy()
    assert got.shape == (N, ops.b1_pitch(K)) and ops.b1_pitch(K) == 16 * ((K + 127) // 128)
    bits = (W.float().numpy() >= 0.0).astype(np.uint8)
    want = np.zeros_like(got)
    pk = np.packbits(bits, axis=1, bitorder="little")
    want[:, : pk.shape[1]] = pk
    assert np.array_equal(got, want)                      # padding bits are zero


@pytest.mark.parametrize("mode", ["binary", "binary01"])
@pytest.mark.parametrize("N,K", [(33, 100), (64, 128), (128, 4096), (9, 4500)])
def test_expand_b1_equals_format_c(cuda, N, K, mode):
    W = _w(N, K, 3 * N + K).to(cuda)
    p1 = ops.quantize_pack_b1(W)
    assert torch.equal(ops.expand_b1(p1, K, mode == "binary01"), ops.quantize_pack(W, mode))


@pytest.mark.parametrize("mode", ["binary", "binary01"])
@pytest.mark.parametrize("M", [1, 2, 3, 5, 8])
@pytest.mark.parametrize("N,K", [(48, 100), (37, 259), (512, 2048), (1000, 4096), (130, 11008), (70, 4500)])
def test_gemv_b1_matches_oracle(cuda, M, N, K, mode):
    g = torch.Generator().manual_seed(M * 1009 + N + K)
    W = _w(N, K, M + N + K)
    x = torch.randn(M, K, generator=g).bfloat16()
    scale = torch.rand(N, generator=g) / K ** 0.5
    bias = torch.randn(N, generator=g) * 0.3
    p1 = ops.quantize_pack_b1(W.to(cuda))
    out = ops.binary_matmul(x.to(cuda), p1, K, scale.to(cuda), bias.to(cuda), 100.0, mode == "binary01", out_dtype=torch.float32,
                            force_b1=True)                     # quantiser + tb200_gemv_b1 at every M <= 8
    if M <= 2:                                                # the one-launch form the layers use
        fused = ops.binary_matmul(x.to(cuda), p1, K, scale.to(cuda), bias.to(cuda), 100.0, mode == "binary01", out_dtype=torch.float32)
        assert torch.equal(fused, out)
    codes = ref_int8.codes_of(W.numpy(), mode)
    qo, so, _ = ref_int8.quant_act(x.float().numpy())
    acc_o = ref_int8.int_acc(qo, codes)                   # exact integers
    y_o = ref_int8.fused_out(acc_o, so, scale.numpy(), bias.numpy(), 100.0)
    got = out.cpu().numpy()
    assert np.allclose(got, y_o, rtol=2e-6, atol=1e-6)    # integer sums identical: only the fp32 epilogue's FMA differs
    # tie to the reference: its fake-quant layer on the de-quantised activations
    xd = torch.from_numpy(qo.astype(np.float32) * so[:, None])
    y_ref = om.scaled_linear_forward(xd, W, scale, bias, 100.0, mode).numpy()
    tol = TOL * np.maximum(np.abs(y_ref), np.sqrt((y_ref ** 2).mean()))
    assert np.all(np.abs(got - y_ref) <= tol)
    # and bit-identical to the 2-bit kernels on the same weights (same integer sums, same epilogue)
    p2 = ops.quantize_pack(W.to(cuda), mode)
    y2 = ops.ternary_matmul(x.to(cuda), p2, scale.to(cuda), bias.to(cuda), 100.0, out_dtype=torch.float32)
    assert torch.equal(out, y2)


@pytest.mark.parametrize("cls", ["ScaledBinaryLinear", "ScaledBinary01Linear", "SmoothlyBinarizedLinear"])
def test_layers_one_bit_equals_two_bit(cuda, cls):
    torch.manual_seed(5)
    layer = getattr(tb, cls)(300, 200).to(cuda).eval()
    if hasattr(layer, "scale"):
        layer.scale.data.fill_(0.05)
    ys = {}
    for M in (1, 7, 8, 9, 40):                            # GEMV, GEMV, GEMV, expanded -> 2-bit GEMV / tcgen05 GEMM
        x = torch.randn(M, 300, generator=torch.Generator().manual_seed(M)).bfloat16().to(cuda)
        with torch.no_grad():
            layer.set_pack_bits(2)
            y2 = layer(x)
            layer.set_pack_bits(1)
            assert layer.packed_weights().shape == (200, ops.b1_pitch(300))
            y1 = layer(x)
        assert torch.equal(y1, y2), f"M={M}"
        ys[M] = y1
    assert torch.equal(layer.codes(), layer.set_pack_bits(2).codes())
    layer.set_pack_bits(1)
    # a write through .data (the reference's idiom) re-packs the 1-bit form too
    with torch.no_grad():
        x = torch.randn(2, 300, device=cuda).bfloat16()
        y_before = layer(x).clone()
        layer.weights.data[:] = -layer.weights.data
        assert not torch.equal(layer(x), y_before)
    # frozen one-bit layers round-trip through a state dict
    layer.freeze(release_latent=True)
    other = getattr(tb, cls)(300, 200).to(cuda).eval()
    other.load_state_dict(layer.state_dict())
    assert other.pack_bits == 1
    with torch.no_grad():
        assert torch.equal(other(x), layer(x))
    with pytest.raises(ValueError):
        tb.ScaledTernaryLinear(8, 8).set_pack_bits(1)


@pytest.mark.parametrize("kind", ["binary", "binary01"])
def test_one_bit_layer_vs_reference_golden(cuda, golden, kind):
    """The live reference's recorded forward of the binary layers (float activations) against the 1-bit int8 path."""
    W, x = torch.from_numpy(golden["small_W"]), torch.from_numpy(golden["small_x"])
    scale, bias = torch.from_numpy(golden["small_scale"]), torch.from_numpy(golden["small_bias"])
    want = golden[f"small_y_{kind}"]
    p1 = ops.quantize_pack_b1(W.to(cuda))
    got = ops.binary_matmul(x.to(cuda), p1, W.shape[1], scale.to(cuda), bias.to(cuda), 100.0, kind == "binary01").float().cpu().numpy()
    rms = np.sqrt((want ** 2).mean())
    assert np.abs(got - want).max() <= 0.05 * rms          # int8 activation quantisation error (reported, ~1 % of rms), not a parity gate
