import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from ternary_vllm_b200 import ops
dev = torch.device("cuda", 0)
K = 4096
M = int(os.environ.get("M", "1"))
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
for (N, R) in [(1376, 64), (1376, 512), (2752, 64), (2752, 256), (4096, 67), (5504, 64), (5504, 128), (8192, 64), (11008, 64), (11008, 128), (16384, 64)]:
    packed = torch.randint(0, 255, (R, N, K // 4), dtype=torch.uint8, device=dev)
    packed &= 0xAA  # codes 0 or 2 only -> valid
    scale = torch.full((R, N), 1 / 64.0, device=dev); bias = torch.zeros(R, N, device=dev)
    xs = torch.randn(R, M, K, device=dev, dtype=torch.bfloat16)
    out = torch.empty(R, M, N, device=dev, dtype=torch.bfloat16)
    ws = torch.empty(ops.L.lib().tb200_gemv_stream_workspace(M, K, R), dtype=torch.uint8, device=dev)
    t = timed(lambda: ops.gemv_stream(xs, packed, scale, bias, 100.0, out=out, workspace=ws))
    mb = R * N * K / 4 / 1e6
    print(f"N={N:6d} R={R:4d} M={M}: {t:8.1f} us  {mb:7.1f} MB  {mb / t * 1e-3 * 1e3:7.1f} GB/s... {mb*1e6/ (t*1e-6)/1e12:.2f} TB/s  blocks/CTA={R*((N+15)//16)/148:.1f}", flush=True)
    del packed, scale, bias, xs, out, ws
