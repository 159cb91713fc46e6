import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from ternary_vllm_b200 import ops
dev = torch.device("cuda", 0)
K = 4096
def timed(fn, it=20):
    st = torch.cuda.Stream(dev)
    fn(); torch.cuda.synchronize()
    with torch.cuda.stream(st):
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=st): fn()
    torch.cuda.synchronize()
    for _ in range(3): gr.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): gr.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it * 1e3
res = []
for (N, R) in [(4096, 67), (11008, 64)]:
    packed = torch.randint(0, 255, (R, N, K // 4), dtype=torch.uint8, device=dev) & 0xAA
    scale = torch.full((R, N), 1 / 64.0, device=dev); bias = torch.zeros(R, N, device=dev)
    for M in (16, 32):
        xs = torch.randn(R, M, K, device=dev, dtype=torch.bfloat16)
        out = torch.empty(R, M, N, device=dev, dtype=torch.bfloat16)
        ws = torch.empty(ops.L.lib().tb200_gemv_stream_workspace(M, K, R), dtype=torch.uint8, device=dev)
        t = timed(lambda: ops.gemv_stream(xs, packed, scale, bias, 100.0, out=out, workspace=ws))
        res.append(f"N={N} M={M}: {t:.1f}us {R*N*K/4/t/1e6:.2f}TB/s")
print("GEMM_DBG", os.environ.get("TB200_GEMM_DBG"), " | ".join(res), flush=True)
