"""tb200_linear_forward_static: the allocation-free variant of tb200_linear_forward (no counterpart in the reference; added
for callers that cannot call cudaMalloc).  Its scratch is a file-scope device array
of TB200_STATIC_SCRATCH_BYTES inside the library.  Called through the C-ABI and compared with the existing entry point
(bit-identical: same kernels, only the scratch address differs) and with the CPU oracle, on two calls in a row with
different inputs (the array keeps the previous call's contents), on every launch form (fused GEMV, quantiser + GEMV,
quantiser + tcgen05 GEMM).  Collected last on purpose (pytest -x): it passes alone in a fresh process and after the rest."""
import numpy as np
import pytest
import torch

from oracle import ref_int8
from ternary_vllm_b200 import _lib as L, ops

pytestmark = pytest.mark.gpu


def _case(M, K, N, seed, dtype, cuda):
    g = torch.Generator().manual_seed(seed)
    W = torch.randn(N, K, generator=g).clip(-1, 1)
    scale, bias = torch.rand(N, generator=g) / K ** 0.5, torch.randn(N, generator=g)
    x = torch.randn(M, K, generator=g).to(dtype)
    return W, x, scale, bias, ops.quantize_pack(W.to(cuda), "ternary")


def _oracle(W, x, scale, bias, clip, dtype):
    q, s, _ = ref_int8.quant_act(x.float().numpy())
    y = ref_int8.fused_out(ref_int8.int_acc(q, ref_int8.codes_of(W.numpy())), s, scale.numpy(), bias.numpy(), clip)
    return torch.from_numpy(y).to(dtype).float().numpy()


def _static(x, packed, sc, bi, clip, N, stream):
    M, K = x.shape
    out = torch.full((M, N), float("nan"), dtype=x.dtype, device=x.device)
    L.check(L.lib().tb200_linear_forward_static(x.data_ptr(), ops._dt(x), M, K, x.stride(0), packed.data_ptr(), packed.stride(0), N,
                                                sc.data_ptr(), bi.data_ptr(), clip, out.data_ptr(), ops._dt(x), out.stride(0),
                                                stream.cuda_stream), "linear_forward_static")
    return out


def _with_workspace(x, packed, sc, bi, clip, N, stream):
    M, K = x.shape
    out = torch.full((M, N), float("nan"), dtype=x.dtype, device=x.device)
    ws = torch.empty(L.lib().tb200_linear_workspace(M, K), dtype=torch.uint8, device=x.device)
    L.check(L.lib().tb200_linear_forward(x.data_ptr(), ops._dt(x), M, K, x.stride(0), packed.data_ptr(), packed.stride(0), N,
                                         sc.data_ptr(), bi.data_ptr(), clip, out.data_ptr(), ops._dt(x), out.stride(0),
                                         ws.data_ptr(), stream.cuda_stream), "linear_forward")
    return out


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("M,K,N", [(1, 1024, 512), (8, 4096, 1000), (12, 1024, 300), (32, 2048, 640), (300, 1024, 384), (729, 4304, 256)])
def test_static_scratch_matches_workspace_call(cuda, M, K, N, dtype):
    stream = torch.cuda.current_stream(cuda)
    tol = {torch.bfloat16: 2.0 ** -7, torch.float32: 2e-6}[dtype]
    for call in range(2):                       # two calls in a row, different inputs: the shared array holds call 0's rows
        W, x, scale, bias, packed = _case(M, K, N, 1000 * call + M + K + N, dtype, cuda)
        xd, sc, bi = x.to(cuda), scale.to(cuda), bias.to(cuda)
        want = _with_workspace(xd, packed, sc, bi, 100.0, N, stream)
        # everything the call needs exists now: from here to the synchronize the caching allocator must not be asked for memory
        torch.cuda.synchronize(cuda)
        out = torch.full((M, N), float("nan"), dtype=dtype, device=cuda)
        before = torch.cuda.memory_stats(cuda).get("allocation.all.allocated", 0)
        L.check(L.lib().tb200_linear_forward_static(xd.data_ptr(), ops._dt(xd), M, K, xd.stride(0), packed.data_ptr(), packed.stride(0), N,
                                                    sc.data_ptr(), bi.data_ptr(), 100.0, out.data_ptr(), ops._dt(xd), out.stride(0),
                                                    stream.cuda_stream), "linear_forward_static")
        torch.cuda.synchronize(cuda)
        assert torch.cuda.memory_stats(cuda).get("allocation.all.allocated", 0) == before      # torch's allocator was not asked
        assert not torch.isnan(out).any()
        assert torch.equal(out, want), f"call {call}: static-scratch call differs from tb200_linear_forward"
        ref = _oracle(W, x, scale, bias, 100.0, dtype)
        got = out.float().cpu().numpy()
        assert np.all(np.abs(got - ref) <= np.abs(ref) * tol + 1e-6), f"call {call}: {np.abs(got - ref).max()}"


def test_static_scratch_alternating_shapes(cuda):
    """Calls of different shapes and launch forms back to back on one stream: a later call must not see the earlier one's rows."""
    stream = torch.cuda.current_stream(cuda)
    shapes = [(300, 1024, 384), (12, 2048, 200), (64, 512, 128), (300, 1024, 384), (3, 4096, 96)]
    cases = [_case(M, K, N, 77 + i, torch.bfloat16, cuda) for i, (M, K, N) in enumerate(shapes)]
    outs = [_static(x.to(cuda), pk, sc.to(cuda), bi.to(cuda), 100.0, W.shape[0], stream) for W, x, sc, bi, pk in cases]
    for (W, x, sc, bi, pk), out in zip(cases, outs):
        assert torch.equal(out, _with_workspace(x.to(cuda), pk, sc.to(cuda), bi.to(cuda), 100.0, W.shape[0], stream))


def test_static_scratch_under_graph_capture(cuda):
    """The use the entry point exists for: captured into a CUDA graph with no workspace to allocate, replayed twice."""
    M, K, N = 40, 1024, 256
    W, x, scale, bias, packed = _case(M, K, N, 5, torch.bfloat16, cuda)
    xd, sc, bi = x.to(cuda), scale.to(cuda), bias.to(cuda)
    out = torch.zeros(M, N, dtype=torch.bfloat16, device=cuda)
    side = torch.cuda.Stream(cuda)
    args = lambda s: (xd.data_ptr(), ops._dt(xd), M, K, K, packed.data_ptr(), packed.stride(0), N, sc.data_ptr(), bi.data_ptr(), 100.0,
                      out.data_ptr(), ops._dt(xd), N, s.cuda_stream)
    with torch.cuda.stream(side):
        L.check(L.lib().tb200_linear_forward_static(*args(side)), "warm-up")          # set-up (symbol address, tensor maps) outside the capture
    side.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side):
        L.check(L.lib().tb200_linear_forward_static(*args(torch.cuda.current_stream(cuda))), "captured")
    want = _with_workspace(xd, packed, sc, bi, 100.0, N, torch.cuda.current_stream(cuda))
    for _ in range(2):
        out.zero_()
        torch.cuda.synchronize(cuda)
        graph.replay()
        torch.cuda.synchronize(cuda)
        assert torch.equal(out, want)


def test_static_scratch_rejects_oversize(cuda):
    lib = L.lib()
    cap = L.STATIC_SCRATCH_BYTES
    M, K = 4096, 4096
    assert lib.tb200_linear_workspace(M, K) > cap
    rc = lib.tb200_linear_forward_static(None, L.TB200_BF16, M, K, K, None, 1024, 16, None, None, 100.0, None, L.TB200_BF16, 16, None)
    assert rc == -1                               # TB200_EINVAL
    msg = lib.tb200_last_error().decode()
    assert "TB200_STATIC_SCRATCH_BYTES" in msg and "4096" in msg
    # the largest token count that fits at K = 4096 is accepted by the size check (and computes)
    M_fit = max(m for m in (1024, 1536, 2000, 2040) if lib.tb200_linear_workspace(m, K) <= cap)
    W, x, scale, bias, packed = _case(M_fit, K, 128, 9, torch.bfloat16, cuda)
    st = torch.cuda.current_stream(cuda)
    got = _static(x.to(cuda), packed, scale.to(cuda), bias.to(cuda), 100.0, 128, st)
    assert torch.equal(got, _with_workspace(x.to(cuda), packed, scale.to(cuda), bias.to(cuda), 100.0, 128, st))
