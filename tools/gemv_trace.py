"""Per-CTA timeline of one decode GEMV launch (clock64 stamps)."""
import math, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
from ternary_vllm_b200 import ops, _lib as L
dev = torch.device("cuda:0")
K, N, M = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
PDL = len(sys.argv) > 4 and sys.argv[4] == "pdl"
g = torch.Generator(device=dev).manual_seed(0)
R = 30
packed = [ops.quantize_pack(torch.randn(N, K, generator=g, device=dev).clip_(-1, 1), "ternary") for _ in range(R)]
x = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
trace = torch.zeros(148 * 16, dtype=torch.int64, device=dev)
if PDL:     # a chain of R kernels under programmatic dependent launch inside a CUDA graph; the last one is traced
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for r in range(R): ops.gemv(x, packed[r], None, None, 100.0, pdl=True, out=out)
        s.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=s):
            for r in range(R - 1): ops.gemv(x, packed[r], None, None, 100.0, pdl=True, out=out)
            L.lib().tb200_debug_set_trace(trace.data_ptr())
            ops.gemv(x, packed[R - 1], None, None, 100.0, pdl=True, out=out)
            L.lib().tb200_debug_set_trace(None)
        for _ in range(3): gr.replay()
    torch.cuda.synchronize()
else:
    for r in range(R - 1):
        ops.gemv(x, packed[r], None, None, 100.0)
    torch.cuda.synchronize()
    L.lib().tb200_debug_set_trace(trace.data_ptr())
    ops.gemv(x, packed[R - 1], None, None, 100.0)
    torch.cuda.synchronize()
L.lib().tb200_debug_set_trace(None)
t = trace.cpu().numpy().reshape(148, 16).astype(np.float64)
t0 = t[:, 0:1]
rel = t - t0
names = ["start", "prologue_done"] + [f"tile{i}_{k}" for i in range(4) for k in ("arr", "done")] + ["bar_init", "x_arrived", "amax_bar", "quantised", "wait_return", "end"]
print("global start spread (clk):", t[:, 0].max() - t[:, 0].min())
for i, n in enumerate(names):
    col = rel[:, i][t[:, i] > 0]
    if len(col): print(f"{n:14s} median {np.median(col):8.0f}  min {col.min():8.0f}  max {col.max():8.0f}  (n={len(col)})")
print("kernel span (clk): ", t[:, 15].max() - t[:, 0].min())
