"""GPU parity, tcgen05 kind::i8 GEMM (prefill shapes): accumulators bit-exact vs the oracle."""
import numpy as np
import pytest
import torch

from oracle import ref_int8, ref_modules as om
from ternary_vllm_b200 import ops

pytestmark = pytest.mark.gpu


def run_case(cuda, M, N, K, seed, out_dtype=torch.bfloat16, scale_mode="sqrt", mode="ternary", bias_on=True):
    g = torch.Generator().manual_seed(seed)
    W = torch.randn(N, K, generator=g).clip(-1, 1)
    x = torch.randn(M, K, generator=g).bfloat16()
    bias = torch.randn(N, generator=g) if bias_on else None
    scale = torch.full((N,), 1.0 / K ** 0.5) if scale_mode == "sqrt" else torch.ones(N)
    packed = ops.quantize_pack(W.to(cuda), mode)
    q, s, sq = ops.quant_act_i8(x.to(cuda), kperm=True)
    out, acc = ops.gemm_i8(q, s, sq, K, packed, scale.to(cuda), None if bias is None else bias.to(cuda), 100.0,
                           out_dtype=out_dtype, debug=True)
    torch.cuda.synchronize()
    codes = ref_int8.codes_of(W.numpy(), mode)
    qo, so, _ = ref_int8.quant_act(x.float().numpy())
    acc_o = ref_int8.int_acc(qo, codes)
    got_acc = acc.cpu().numpy().astype(np.int64)
    bad = np.argwhere(got_acc != acc_o)
    assert bad.shape[0] == 0, f"{bad.shape[0]} accumulators differ, first at {bad[:5].tolist()}: got {got_acc[tuple(bad[0])]} want {acc_o[tuple(bad[0])]}"
    y_o = ref_int8.fused_out(acc_o, so, scale.numpy(), None if bias is None else bias.numpy(), 100.0)
    got = out.float().cpu().numpy()
    if out_dtype == torch.float32:
        assert np.allclose(got, y_o, rtol=2e-6, atol=1e-6)
    else:
        want = torch.from_numpy(y_o).to(out_dtype).float().numpy()
        ulp = np.abs(want) * (2.0 ** -7 if out_dtype == torch.bfloat16 else 2.0 ** -10) + 1e-6
        assert np.all(np.abs(got - want) <= ulp)
    return W, x, scale, bias, qo, so, y_o


@pytest.mark.parametrize("M,N,K", [
    (256, 128, 128),        # one tile, one k-block
    (256, 128, 512),        # pipeline wraps (4 stages)
    (256, 256, 1024),       # two weight tiles, phase flips
    (512, 384, 256),        # several tiles per CTA -> TMEM double buffering
    (300, 200, 320),        # ragged M, N, K (K % 128 != 0)
    (257, 129, 100),        # K < 128, K % 16 != 0
    (33, 130, 256), (64, 128, 256), (100, 64, 384), (16, 128, 128),   # narrow token tiles (BT 64/128/16)
])
def test_gemm_parity_small(cuda, M, N, K):
    run_case(cuda, M, N, K, seed=M + 3 * N + 7 * K)


def test_gemm_many_tiles_persistent(cuda):
    # 8 x 40 = 320 tiles on 148 CTAs: every CTA walks 2-3 tiles, ragged tail
    run_case(cuda, 2048, 5000, 512, seed=1)


@pytest.mark.parametrize("out_dtype", [torch.float32, torch.float16])
def test_gemm_out_dtypes_and_clip(cuda, out_dtype):
    W, x, scale, bias, qo, so, y_o = run_case(cuda, 384, 256, 2048, seed=2, out_dtype=out_dtype, scale_mode="one")
    assert (np.abs(y_o) == 100.0).mean() > 0.001


@pytest.mark.parametrize("mode", ["binary", "binary01"])
def test_gemm_binary_modes(cuda, mode):
    run_case(cuda, 256, 128, 256, seed=3, mode=mode, bias_on=False)


def test_gemm_vs_reference_layer_on_dequantised_activations(cuda):
    W, x, scale, bias, qo, so, y_o = run_case(cuda, 320, 192, 1024, seed=4, out_dtype=torch.float32)
    xd = torch.from_numpy(qo.astype(np.float32) * so[:, None])
    y_ref = om.scaled_linear_forward(xd, W, scale, bias, 100.0, "ternary").numpy()
    tol = 1e-3 * np.maximum(np.abs(y_ref), np.sqrt((y_ref ** 2).mean()))
    assert np.all(np.abs(y_o - y_ref) <= tol)


def exact_acc_cpu(x_bf16: torch.Tensor, W: torch.Tensor, mode="ternary"):
    """Exact integer accumulators of the WHOLE problem on the host cores: int8 codes and {-1,0,1} weights as float32
    through BLAS.  Every partial sum is an integer of magnitude <= 127*K < 2^24, so fp32 accumulation is exact in any
    order (K <= 131072)."""
    qo, so, sqo = ref_int8.quant_act(x_bf16.float().numpy())
    codes = torch.from_numpy(ref_int8.codes_of(W.numpy(), mode)).float()
    assert 127 * W.shape[1] < 2 ** 24
    acc = (torch.from_numpy(qo).float() @ codes.T).to(torch.int32)
    return qo, so, sqo, acc


def test_gemm_full_size_exact(cuda):
    """BASELINE configs[2] at full size (M=4096, K=4096, N=11008): ALL 45 M accumulators against the exact CPU product,
    and all outputs against the oracle epilogue (one bf16 rounding)."""
    M, N, K = 4096, 11008, 4096
    g = torch.Generator().manual_seed(11)
    W = torch.randn(N, K, generator=g).clip(-1, 1)
    x = torch.randn(M, K, generator=g).bfloat16()
    scale, bias = torch.full((N,), 1.0 / K ** 0.5), torch.randn(N, generator=g)
    packed = ops.quantize_pack(W.to(cuda), "ternary")
    q, s, sq = ops.quant_act_i8(x.to(cuda), kperm=True)
    out, acc = ops.gemm_i8(q, s, sq, K, packed, scale.to(cuda), bias.to(cuda), 100.0, out_dtype=torch.bfloat16, debug=True)
    qo, so, sqo, acc_o = exact_acc_cpu(x, W)
    assert np.array_equal(s.cpu().numpy(), so) and np.array_equal(sq.cpu().numpy(), sqo)
    assert torch.equal(acc.cpu(), acc_o), f"{(acc.cpu() != acc_o).sum().item()} of {acc_o.numel()} accumulators differ"
    y_o = ref_int8.fused_out(acc_o.numpy(), so, scale.numpy(), bias.numpy(), 100.0)
    want = torch.from_numpy(y_o).bfloat16().float()
    assert torch.all((out.float().cpu() - want).abs() <= want.abs() * 2.0 ** -7 + 1e-6)
    # the layer call (quantiser + GEMM, what tb200_linear_forward runs) gives the same bits
    assert torch.equal(ops.ternary_matmul(x.to(cuda), packed, scale.to(cuda), bias.to(cuda), 100.0), out)


def test_gemm_full_size_properties(cuda):
    """BASELINE configs[2] at full size (M=4096, K=4096, N=11008): too big for the CPU oracle in
    seconds, so check size-independent properties: (i) a random sample of accumulators against the
    oracle, (ii) agreement with the independent dp4a GEMV kernel on 32 token rows, (iii) linearity
    in the weights: acc(W) on x equals -acc(-W)."""
    M, N, K = 4096, 11008, 4096
    g = torch.Generator().manual_seed(11)
    W = torch.randn(N, K, generator=g).clip(-1, 1)
    x = torch.randn(M, K, generator=g).bfloat16()
    Wc, xc = W.to(cuda), x.to(cuda)
    packed = ops.quantize_pack(Wc, "ternary")
    q, s, sq = ops.quant_act_i8(xc, kperm=True)
    out, acc = ops.gemm_i8(q, s, sq, K, packed, None, None, 1e9, out_dtype=torch.float32, debug=True)
    rows = torch.randint(0, M, (24,), generator=g)
    cols = torch.randint(0, N, (64,), generator=g)
    qo, so, _ = ref_int8.quant_act(x[rows].float().numpy())
    codes = ref_int8.codes_of(W[cols].numpy())
    want = ref_int8.int_acc(qo, codes)
    assert np.array_equal(acc[rows.to(cuda)][:, cols.to(cuda)].cpu().numpy().astype(np.int64), want)
    _, acc_v, _, _ = ops.gemv(xc[:32].contiguous(), packed, None, None, 1e9, debug=True)
    assert torch.equal(acc_v, acc[:32])
    packed_neg = ops.quantize_pack(-Wc, "ternary")
    _, acc_n = ops.gemm_i8(q, s, sq, K, packed_neg, None, None, 1e9, out_dtype=torch.float32, debug=True)
    assert torch.equal(acc_n, -acc)
    # the last tile column / row were written
    assert torch.isfinite(out[-1]).all() and torch.isfinite(out[:, -1]).all()


@pytest.mark.parametrize("K,N", [(1024, 3072), (1024, 1024), (1024, 4096), (4096, 1024)])
def test_vit_l14_shapes(cuda, K, N):
    """BASELINE configs[3]: ViT-L/14 linears at 32 images x 257 tokens (M = 8224 = 42 x 192 + 160: ragged token tile).
    ALL accumulators against the exact CPU product, plus agreement with the GEMV kernel on the last 32 tokens."""
    M = 32 * 257
    g = torch.Generator().manual_seed(K + N)
    W = torch.randn(N, K, generator=g).clip(-1, 1)
    x = torch.randn(M, K, generator=g).bfloat16()
    packed = ops.quantize_pack(W.to(cuda), "ternary")
    q, s, sq = ops.quant_act_i8(x.to(cuda), kperm=True)
    out, acc = ops.gemm_i8(q, s, sq, K, packed, None, None, 1e9, out_dtype=torch.float32, debug=True)
    qo, so, sqo, acc_o = exact_acc_cpu(x, W)
    assert np.array_equal(s.cpu().numpy(), so) and np.array_equal(sq.cpu().numpy(), sqo)
    assert torch.equal(acc.cpu(), acc_o), f"{(acc.cpu() != acc_o).sum().item()} of {acc_o.numel()} accumulators differ"
    assert np.allclose(out.cpu().numpy(), ref_int8.fused_out(acc_o.numpy(), so, None, None, 1e9), rtol=2e-6, atol=1e-6)
    _, acc_v, _, _ = ops.gemv(x[-32:].to(cuda).contiguous(), packed, None, None, 1e9, debug=True)
    assert torch.equal(acc_v, acc[-32:])
    y = ops.ternary_matmul(x.to(cuda).reshape(32, 257, K), packed, None, None, 1e9)
    assert y.shape == (32, 257, N) and torch.isfinite(y).all()


@pytest.mark.parametrize("M,N,K,world", [(300, 1000, 1024, 2), (513, 704, 4096, 3), (40, 256, 512, 4)])
def test_prefill_tp_peer_stores_single_device(cuda, M, N, K, world):
    """tb200_linear_forward_tp for MORE than 32 tokens on ONE device: every 'rank' runs quantiser + tcgen05 GEMM on its row
    block and its epilogue stores the tile into `world` local stand-ins for the peer buffers (TMA tile stores, ragged last
    shard, column offsets); every buffer must equal the unsharded tb200_linear_forward output bit for bit."""
    import ctypes

    from ternary_vllm_b200 import _lib as L
    from ternary_vllm_b200.parallel import shard_rows, shard_size

    g = torch.Generator().manual_seed(M + N + K)
    W = torch.randn(N, K, generator=g).clip(-1, 1).to(cuda)
    x = torch.randn(M, K, generator=g).bfloat16().to(cuda)
    scale = (torch.rand(N, generator=g) / K ** 0.5).to(cuda)
    bias = torch.randn(N, generator=g).to(cuda)
    packed = ops.quantize_pack(W, "ternary")
    want = ops.ternary_matmul(x, packed, scale, bias, 100.0)
    n_shard = shard_size(N, world)
    ld = world * n_shard
    bufs = [torch.full((M, ld), float("nan"), dtype=torch.bfloat16, device=cuda) for _ in range(world)]
    ws = torch.empty(L.lib().tb200_linear_workspace(M, K), dtype=torch.uint8, device=cuda)
    for rank in range(world):
        a, b = shard_rows(N, world, rank)
        if b <= a:
            continue
        ptrs = (ctypes.c_void_p * world)(*[buf.data_ptr() + rank * n_shard * 2 for buf in bufs])
        pk, sc, bi = packed[a:b].contiguous(), scale[a:b].contiguous(), bias[a:b].contiguous()
        L.check(L.lib().tb200_linear_forward_tp(x.data_ptr(), L.TB200_BF16, M, K, K, pk.data_ptr(), pk.stride(0), b - a,
                                                sc.data_ptr(), bi.data_ptr(), 100.0, ptrs, world, L.TB200_BF16, ld,
                                                ws.data_ptr(), torch.cuda.current_stream().cuda_stream), "linear_forward_tp")
    torch.cuda.synchronize()
    for buf in bufs:
        assert torch.equal(buf[:, :N], want)
