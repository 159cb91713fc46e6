"""GPU parity of the HOST-BUFFER entry points (the calls bench.py's `e2e` times): tb200_linear_forward_host and
tb200_gemv_stream_host take pinned host activations, copy them in, run the layer(s) and copy the outputs back.
Outputs are compared with the CPU oracle exactly as the device-buffer tests do (one rounding to the output type)."""
import numpy as np
import pytest
import torch

from oracle import ref_int8
from ternary_vllm_b200 import _lib as L, ops

pytestmark = pytest.mark.gpu


def oracle_out(W, x, scale, bias, clip, out_dtype):
    q, s, _ = ref_int8.quant_act(x.float().numpy())
    y = ref_int8.fused_out(ref_int8.int_acc(q, ref_int8.codes_of(W.numpy())), s, scale.numpy(), bias.numpy(), clip)
    return torch.from_numpy(y).to(out_dtype).float().numpy() if out_dtype != torch.float32 else y


def assert_close(got, want, dtype):
    ulp = np.abs(want) * {torch.bfloat16: 2.0 ** -7, torch.float16: 2.0 ** -10, torch.float32: 2e-6}[dtype] + 1e-6
    assert np.all(np.abs(got - want) <= ulp), np.abs(got - want).max()


@pytest.mark.parametrize("M,K,N", [(1, 1024, 512), (8, 4096, 1000), (12, 1024, 300), (32, 2048, 640), (300, 1024, 384)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_linear_forward_host(cuda, M, K, N, dtype):
    g = torch.Generator().manual_seed(M + K + N)
    W = torch.randn(N, K, generator=g).clip(-1, 1)
    scale, bias = torch.full((N,), 1.0 / K ** 0.5), torch.randn(N, generator=g)
    x = torch.randn(M, K, generator=g).to(dtype)
    packed = ops.quantize_pack(W.to(cuda), "ternary")
    hx, hy = x.clone().pin_memory(), torch.full((M, N), float("nan"), dtype=dtype).pin_memory()
    dx, dy = torch.empty(M, K, dtype=dtype, device=cuda), torch.empty(M, N, dtype=dtype, device=cuda)
    ws = torch.empty(L.lib().tb200_linear_workspace(M, K), dtype=torch.uint8, device=cuda)
    sc, bi = scale.to(cuda), bias.to(cuda)
    st = torch.cuda.Stream(cuda)
    L.check(L.lib().tb200_linear_forward_host(hx.data_ptr(), ops._dt(x), M, K, packed.data_ptr(), packed.stride(0), N,
                                              sc.data_ptr(), bi.data_ptr(), 100.0, hy.data_ptr(), ops._dt(x),
                                              dx.data_ptr(), dy.data_ptr(), ws.data_ptr(), st.cuda_stream), "linear_forward_host")
    st.synchronize()
    assert_close(hy.float().numpy(), oracle_out(W, x, scale, bias, 100.0, dtype), dtype)
    # and the same bits as the device-buffer call
    assert torch.equal(hy, ops.ternary_matmul(x.to(cuda), packed, sc, bi, 100.0).cpu())


@pytest.mark.parametrize("M", [1, 8, 9, 32])
@pytest.mark.parametrize("K,N,R", [(1024, 304, 3), (4096, 1000, 2)])
def test_gemv_stream_host(cuda, M, K, N, R):
    g = torch.Generator().manual_seed(M * 7 + K + N)
    Ws = [torch.randn(N, K, generator=g).clip(-1, 1) for _ in range(R)]
    scale = torch.rand(R, N, generator=g) / K ** 0.5
    bias = torch.randn(R, N, generator=g)
    xs = torch.randn(R, M, K, generator=g).bfloat16()
    packed = torch.stack([ops.quantize_pack(W.to(cuda), "ternary") for W in Ws])
    hx, hy = xs.clone().pin_memory(), torch.full((R, M, N), float("nan"), dtype=torch.bfloat16).pin_memory()
    dx, dy = torch.empty_like(xs, device=cuda), torch.empty(R, M, N, dtype=torch.bfloat16, device=cuda)
    ws = torch.empty(L.lib().tb200_gemv_stream_workspace(M, K, R), dtype=torch.uint8, device=cuda)
    sc, bi = scale.to(cuda), bias.to(cuda)
    st = torch.cuda.Stream(cuda)
    L.check(L.lib().tb200_gemv_stream_host(hx.data_ptr(), L.TB200_BF16, M, K, packed.data_ptr(), packed.stride(1),
                                           packed.stride(0), N, sc.data_ptr(), bi.data_ptr(), N, 100.0, hy.data_ptr(),
                                           L.TB200_BF16, R, dx.data_ptr(), dy.data_ptr(), ws.data_ptr(), st.cuda_stream),
            "gemv_stream_host")
    st.synchronize()
    for r in range(R):
        assert_close(hy[r].float().numpy(), oracle_out(Ws[r], xs[r], scale[r], bias[r], 100.0, torch.bfloat16), torch.bfloat16)
    assert torch.equal(hy, ops.gemv_stream(xs.to(cuda), packed, sc, bi, 100.0).cpu())


def test_host_entry_rejects_null(cuda):
    rc = L.lib().tb200_linear_forward_host(None, L.TB200_BF16, 1, 64, None, 16, 4, None, None, 100.0, None, L.TB200_BF16,
                                           None, None, None, None)
    assert rc == -1 and b"null" in L.lib().tb200_last_error()
