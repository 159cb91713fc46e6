"""Cost of the PEER (tensor-parallel epilogue) instantiation of the stream GEMV on ONE GPU: the same ring through tb200_gemv_stream
and through tb200_gemv_stream_tp with a single LOCAL destination (no NVLink involved)."""
import ctypes, math, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ternary_vllm_b200 import ops, _lib as L

dev = torch.device("cuda:0")
K, N, R = 4096, int(sys.argv[1]) if len(sys.argv) > 1 else 11008, int(sys.argv[2]) if len(sys.argv) > 2 else 64
g = torch.Generator(device=dev).manual_seed(0)
ld = ops.packed_pitch(K)
packed = torch.empty(R, N, ld, dtype=torch.uint8, device=dev)
for r in range(R):
    packed[r] = ops.quantize_pack(torch.randn(N, K, generator=g, device=dev).clip_(-1, 1), "ternary")
scale = torch.full((R, N), 1 / math.sqrt(K), device=dev); bias = torch.randn(R, N, generator=g, device=dev)
st = torch.cuda.Stream(dev)
for M in (1, 8):
    xs = torch.randn(R, M, K, generator=g, device=dev, dtype=torch.bfloat16)
    out = torch.empty(R, M, N, device=dev, dtype=torch.bfloat16)
    out2 = torch.empty(R, M, N, device=dev, dtype=torch.bfloat16)
    ws = torch.empty(L.lib().tb200_gemv_stream_workspace(M, K, R), dtype=torch.uint8, device=dev)
    ptrs = (ctypes.c_void_p * 1)(out2.data_ptr())
    fns = {"plain": lambda: ops.gemv_stream(xs, packed, scale, bias, 100.0, out=out, workspace=ws),
           "peer(1 local)": lambda: ops.gemv_stream_tp(xs, packed, scale, bias, 100.0, ptrs, 1, torch.bfloat16, N, M * N, workspace=ws)}
    with torch.cuda.stream(st):
        for name, fn in fns.items():
            fn(); st.synchronize()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=st):
                fn()
            for _ in range(3): gr.replay()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            st.synchronize(); e0.record(st)
            for _ in range(20): gr.replay()
            e1.record(st); st.synchronize()
            print(f"N={N} R={R} M={M} {name}: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us per call", flush=True)
    assert os.environ.get("TB200_STREAM_DBG") or torch.equal(out, out2)
