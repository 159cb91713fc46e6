"""tcgen05 GEMM timing probe: M tokens x (K -> N), both weight paths."""
import math, os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ternary_vllm_b200 import ops
dev = torch.device("cuda:0")
shapes = [(4096, 4096, 11008)] if len(sys.argv) < 2 else [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]]
for (M, K, N) in shapes:
    g = torch.Generator(device=dev).manual_seed(7)
    packed = ops.quantize_pack(torch.randn(N, K, generator=g, device=dev).clip_(-1, 1), "ternary")
    x = torch.randn(M, K, generator=g, device=dev, dtype=torch.bfloat16)
    scale = torch.full((N,), 1 / math.sqrt(K), device=dev); bias = torch.randn(N, generator=g, device=dev)
    q, s, sq = ops.quant_act_i8(x, kperm=True)
    flush = torch.empty(int(256e6), dtype=torch.uint8, device=dev)
    for _ in range(3): ops.gemm_i8(q, s, sq, K, packed, scale, bias, 100.0)
    ts = []
    for _ in range(10):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ops.gemm_i8(q, s, sq, K, packed, scale, bias, 100.0); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    t = statistics.median(ts) * 1e-3
    print(f"path={os.environ.get('TB200_GEMM_PATH','ts')} M={M} K={K} N={N}: {t*1e6:8.1f} us  {2*M*N*K/t/1e12:8.1f} TOPS  ({2*M*N*K/t/4.5e15:.3f} of 4.5 POPS)")
    # activation quantisation alone (CUDA graph of 20 launches: the Python call overhead exceeds the kernel)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        ops.quant_act_i8(x, kperm=True); st.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=st):
            for _ in range(20): ops.quant_act_i8(x, kperm=True)
        gr.replay(); st.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st); gr.replay(); e1.record(st); st.synchronize()
    tq = e0.elapsed_time(e1) * 1e-3 / 20
    print(f"   quant_act_i8 M={M} K={K}: {tq*1e6:8.1f} us  {(M*K*3)/tq/1e9:8.1f} GB/s")
