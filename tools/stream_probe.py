"""Device time of ONE tb200_gemv_stream call over a ring of layers, per token count (graph replay, ring larger than L2).
usage: python tools/stream_probe.py [N] [R] [M ...]      (TB200_LIB / TB200_GEMM_DBG / TB200_STREAM_DBG select experiments)"""
import math, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ternary_vllm_b200 import ops, _lib as L

dev = torch.device("cuda:0")
K = 4096
N = int(sys.argv[1]) if len(sys.argv) > 1 else 11008
R = int(sys.argv[2]) if len(sys.argv) > 2 else 64
Ms = [int(v) for v in sys.argv[3:]] or [32]
g = torch.Generator(device=dev).manual_seed(0)
packed = torch.empty(R, N, ops.packed_pitch(K), dtype=torch.uint8, device=dev)
for r in range(R):
    packed[r] = ops.quantize_pack(torch.randn(N, K, generator=g, device=dev).clip_(-1, 1), "ternary")
scale = torch.full((R, N), 1 / math.sqrt(K), device=dev); bias = torch.randn(R, N, generator=g, device=dev)
st = torch.cuda.Stream(dev)
for M in Ms:
    xs = torch.randn(R, M, K, generator=g, device=dev, dtype=torch.bfloat16)
    out = torch.empty(R, M, N, device=dev, dtype=torch.bfloat16)
    ws = torch.empty(L.lib().tb200_gemv_stream_workspace(M, K, R), dtype=torch.uint8, device=dev)
    with torch.cuda.stream(st):
        ops.gemv_stream(xs, packed, scale, bias, 100.0, out=out, workspace=ws); st.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=st):
            ops.gemv_stream(xs, packed, scale, bias, 100.0, out=out, workspace=ws)
        for _ in range(3): gr.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        st.synchronize(); e0.record(st)
        for _ in range(20): gr.replay()
        e1.record(st); st.synchronize()
    us = e0.elapsed_time(e1) / 20 * 1e3
    by = R * (N * K / 4 + M * K * 2 + M * N * 2 + N * 8)
    print(f"N={N} R={R} M={M} dbg={os.environ.get('TB200_GEMM_DBG', '0')}: {us:.1f} us per call, {by / us / 1e3:.0f} GB/s", flush=True)
