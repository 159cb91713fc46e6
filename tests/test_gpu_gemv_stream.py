"""GPU parity, streaming decode GEMV (tb200_gemv_stream): a list of layers in one persistent launch.
Integer accumulators bit-exact vs the oracle; outputs bit-identical to the single-layer kernels, whose
tie to the reference layer is checked in test_gpu_gemv.py."""
import numpy as np
import pytest
import torch

from oracle import ref_int8
from ternary_vllm_b200 import ops

pytestmark = pytest.mark.gpu


def make_layers(R, N, K, M, seed, shared_x=False):
    g = torch.Generator().manual_seed(seed)
    Ws = [torch.randn(N, K, generator=g).clip(-1, 1) for _ in range(R)]
    x = torch.randn((M, K) if shared_x else (R, M, K), generator=g).bfloat16()
    scale = torch.rand(R, N, generator=g) * (2.0 / K ** 0.5)
    bias = torch.randn(R, N, generator=g)
    return Ws, x, scale, bias


def oracle_out(W, x, scale, bias, clip, dtype):
    q, s, _ = ref_int8.quant_act(x.float().numpy())
    acc = ref_int8.int_acc(q, ref_int8.codes_of(W.numpy()))
    y = ref_int8.fused_out(acc, s, scale.numpy(), bias.numpy(), clip)
    return acc, torch.from_numpy(y).to(dtype).float().numpy()


@pytest.mark.parametrize("M", [1, 2, 5, 8, 9, 16, 23, 32])
@pytest.mark.parametrize("R,N,K", [(1, 64, 64), (3, 37, 259), (4, 512, 2048), (2, 1000, 4096), (7, 40, 512),
                                   (3, 2400, 1024), (2, 130, 11008), (2, 100, 5000)])
def test_stream_parity(cuda, M, R, N, K):
    Ws, x, scale, bias = make_layers(R, N, K, M, seed=M * 7919 + N * 31 + K + R)
    packed = torch.stack([ops.quantize_pack(W.to(cuda), "ternary") for W in Ws])
    image_bytes = ((K + 4095) // 4096) * 4096 * min(M, 8) * ((M + 7) // 8)
    try:
        out, acc = ops.gemv_stream(x.to(cuda), packed, scale.to(cuda), bias.to(cuda), 100.0, debug=True)
    except ops.TernaryB200Error as e:        # <= 8 tokens: the activation image must fit shared memory next to a 3-stage ring
        assert "TB200_ENOTSUP" in str(e) and image_bytes > 120 * 1024, str(e)
        return
    torch.cuda.synchronize()
    assert out.shape == (R, M, N) and out.dtype == torch.bfloat16
    for r in range(R):
        acc_o, want = oracle_out(Ws[r], x[r], scale[r], bias[r], 100.0, torch.bfloat16)
        if r == 0:
            assert np.array_equal(acc.cpu().numpy().astype(np.int64), acc_o), "integer accumulators differ"
        got = out[r].float().cpu().numpy()
        assert np.all(np.abs(got - want) <= np.abs(want) * 2.0 ** -7 + 1e-6), f"layer {r}"


@pytest.mark.parametrize("M", [1, 8])
def test_stream_matches_single_layer_kernel(cuda, M):
    """Same bits as tb200_gemv on every layer (identical integer sums and epilogue arithmetic), also when the
    list is longer than the number of SMs' worth of row blocks and with activations shared by all layers."""
    R, N, K = 6, 4096, 4096
    Ws, x, scale, bias = make_layers(R, N, K, M, seed=5 + M, shared_x=True)
    packed = torch.stack([ops.quantize_pack(W.to(cuda), "ternary") for W in Ws])
    xd, sd, bd = x.to(cuda), scale.to(cuda), bias.to(cuda)
    out = ops.gemv_stream(xd, packed, sd, bd, 100.0)
    for r in range(R):
        assert torch.equal(out[r], ops.gemv(xd, packed[r], sd[r], bd[r], 100.0)), f"layer {r}"


@pytest.mark.parametrize("odtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("M", [9, 20, 32])
@pytest.mark.parametrize("R,N,K", [(22, 1800, 4096), (30, 1300, 3000), (40, 1024, 2560), (21, 1920, 3840)])
def test_stream_long_list_resident_activations(cuda, M, R, N, K, odtype):
    """9..32 tokens over a list with several weight tiles per SM and K <= 4096: the tcgen05 kernel keeps a layer's
    activations resident in shared memory and walks contiguous runs of tiles (gemm_ts.cu, TsCfg RES; K = 3840 has an odd
    number of k-blocks and stays on the ring form).  Same bits as the layer-by-layer calls (ring form: too few tiles for
    the resident one), integer accumulators of layer 0 against the oracle, every layer against the oracle's outputs."""
    Ws, x, scale, bias = make_layers(R, N, K, M, seed=M + R + N + K)
    packed = torch.stack([ops.quantize_pack(W.to(cuda), "ternary") for W in Ws])
    xd, sd, bd = x.to(cuda), scale.to(cuda), bias.to(cuda)
    out, acc = ops.gemv_stream(xd, packed, sd, bd, 100.0, out_dtype=odtype, debug=True)
    out2 = ops.gemv_stream(xd, packed, sd, bd, 100.0, out_dtype=odtype)          # a second call: same bits
    torch.cuda.synchronize()
    assert torch.equal(out, out2)
    tol = {torch.bfloat16: 2.0 ** -7, torch.float32: 2e-6}[odtype]
    for r in range(R):
        if M > 16:      # tb200_linear_forward runs the same tcgen05 kernel (ring form) above 16 tokens
            assert torch.equal(out[r], ops.ternary_matmul(xd[r], packed[r], sd[r], bd[r], 100.0, out_dtype=odtype)), f"layer {r}"
        if M <= 16 or r in (0, R // 2, R - 1):
            acc_o, want = oracle_out(Ws[r], x[r], scale[r], bias[r], 100.0, odtype)
            if r == 0:
                assert np.array_equal(acc.cpu().numpy().astype(np.int64), acc_o), "integer accumulators differ"
            got = out[r].float().cpu().numpy()
            assert np.all(np.abs(got - want) <= np.abs(want) * tol + 1e-6), f"layer {r}"


@pytest.mark.parametrize("odtype", [torch.float32, torch.float16, torch.bfloat16])
def test_stream_dtypes_no_scale_bias(cuda, odtype):
    R, N, K, M = 2, 200, 1024, 3
    Ws, x, _, _ = make_layers(R, N, K, M, seed=77)
    xd = x.to(cuda).to(torch.float16 if odtype == torch.float16 else (torch.float32 if odtype == torch.float32 else torch.bfloat16))
    packed = torch.stack([ops.quantize_pack(W.to(cuda), "ternary") for W in Ws])
    out = ops.gemv_stream(xd, packed, None, None, 100.0, out_dtype=odtype)
    for r in range(R):
        q, s, _ = ref_int8.quant_act(xd[r].float().cpu().numpy())
        acc = ref_int8.int_acc(q, ref_int8.codes_of(Ws[r].numpy()))
        y = ref_int8.fused_out(acc, s, None, None, 100.0)
        want = torch.from_numpy(y).to(odtype).float().numpy()
        eps = {torch.float32: 2e-6, torch.float16: 2.0 ** -10, torch.bfloat16: 2.0 ** -7}[odtype]
        assert np.all(np.abs(out[r].float().cpu().numpy() - want) <= np.abs(want) * eps + 1e-5)


def test_stream_repeated_calls_share_workspace(cuda):
    """Back-to-back calls on one stream reuse the blob workspace: the quantiser of call i+1 must wait for call i."""
    R, N, K, M = 3, 512, 2048, 4
    Ws, x, scale, bias = make_layers(R, N, K, M, seed=11)
    packed = torch.stack([ops.quantize_pack(W.to(cuda), "ternary") for W in Ws])
    sd, bd = scale.to(cuda), bias.to(cuda)
    ws = torch.empty(1 << 20, dtype=torch.uint8, device=cuda)
    xs = [torch.randn(R, M, K, generator=torch.Generator().manual_seed(100 + i)).bfloat16().to(cuda) for i in range(6)]
    outs = [ops.gemv_stream(xi, packed, sd, bd, 100.0, workspace=ws) for xi in xs]
    torch.cuda.synchronize()
    for xi, oi in zip(xs, outs):
        assert torch.equal(oi, ops.gemv_stream(xi, packed, sd, bd, 100.0))


@pytest.mark.parametrize("M", [1, 8, 19, 32])
@pytest.mark.parametrize("R,N,K,world", [(3, 400, 1024, 4), (2, 1376, 4096, 8), (2, 100, 5000, 2)])
def test_stream_tp_peer_stores(cuda, M, R, N, K, world):
    """tb200_gemv_stream_tp on ONE device: every 'rank' runs its row block of all layers and stores it into `world`
    local stand-ins for the peer buffers; each buffer must end up bit-identical to the unsharded tb200_gemv_stream
    output (the peer-store epilogues of both the mma.sync and the tcgen05 kernel, column offsets, ragged last shard)."""
    import ctypes

    from ternary_vllm_b200.parallel import shard_rows, shard_size

    Ws, x, scale, bias = make_layers(R, N, K, M, seed=M + N + K + R + world)
    packed = torch.stack([ops.quantize_pack(W.to(cuda), "ternary") for W in Ws])
    xd, sd, bd = x.to(cuda), scale.to(cuda), bias.to(cuda)
    want = ops.gemv_stream(xd, packed, sd, bd, 100.0)
    n_shard = shard_size(N, world)
    ld = world * n_shard
    bufs = [torch.full((R, M, ld), float("nan"), dtype=torch.bfloat16, device=cuda) for _ in range(world)]
    for rank in range(world):
        a, b = shard_rows(N, world, rank)
        if b <= a:
            continue
        ptrs = (ctypes.c_void_p * world)(*[buf.data_ptr() + rank * n_shard * 2 for buf in bufs])
        ops.gemv_stream_tp(xd, packed[:, a:b].contiguous(), sd[:, a:b].contiguous(), bd[:, a:b].contiguous(), 100.0,
                           ptrs, world, torch.bfloat16, ld, M * ld)
    torch.cuda.synchronize()
    for buf in bufs:
        assert torch.equal(buf[:, :, :N], want)
