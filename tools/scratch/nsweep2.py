import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from ternary_vllm_b200 import ops
dev = torch.device("cuda", int(os.environ.get("DEV", "0")))
torch.cuda.set_device(dev)
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
N, R = 5504, 64
g = torch.Generator(device=dev).manual_seed(1)
real = torch.stack([ops.quantize_pack(torch.randn(N, K, generator=g, device=dev).clip_(-1, 1), "ternary") for _ in range(R)])
synth = torch.randint(0, 255, (R, N, K // 4), dtype=torch.uint8, device=dev) & 0xAA
for M in (1, 8):
    for wname, packed in (("real", real), ("synth", synth)):
        for oname, ld, tok in (("compact", N, M), ("strided", 11008, 32)):
            scale = torch.full((R, N), 1 / 64.0, device=dev); bias = torch.zeros(R, N, device=dev)
            xs = torch.randn(R, M, K, device=dev, dtype=torch.bfloat16)
            big = torch.empty(R, tok, ld, device=dev, dtype=torch.bfloat16)
            out = big[:, :M, :N]
            ws = torch.empty(ops.L.lib().tb200_gemv_stream_workspace(M, K, R), dtype=torch.uint8, device=dev)
            t = timed(lambda: ops.gemv_stream(xs, packed, scale, bias, 100.0, out=out, workspace=ws))
            print(f"dev={dev.index} M={M} {wname:6s} {oname:8s}: {t:8.1f} us", flush=True)
