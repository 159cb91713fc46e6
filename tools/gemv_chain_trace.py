"""Cross-kernel timeline (globaltimer, ns) of the last T kernels of a PDL chain of decode GEMVs in a CUDA graph."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
from ternary_vllm_b200 import ops, _lib as L
dev = torch.device("cuda:0")
K, N, M = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
PDL = not (len(sys.argv) > 4 and sys.argv[4] == "nopdl")
g = torch.Generator(device=dev).manual_seed(0)
R, T = 30, 5
packed = [ops.quantize_pack(torch.randn(N, K, generator=g, device=dev).clip_(-1, 1), "ternary") for _ in range(R)]
x = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
trace = torch.zeros(T, 148 * 64, dtype=torch.int64, device=dev)
out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
s = torch.cuda.Stream()
def call(r):
    if M <= 8: ops.gemv(x, packed[r], None, None, 100.0, pdl=PDL, out=out)
    else: ops.ternary_matmul(x, packed[r], None, None, 100.0)
with torch.cuda.stream(s):
    for r in range(R): call(r)
    s.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr, stream=s):
        for r in range(R):
            if r >= R - T: L.lib().tb200_debug_set_trace(trace[r - (R - T)].data_ptr())
            call(r)
        L.lib().tb200_debug_set_trace(None)
    for _ in range(3): gr.replay()
torch.cuda.synchronize()
t = trace.cpu().numpy().reshape(T, 148, 64).astype(np.float64)
names = {0: "start", 1: "wait_return", 3: "image_done", 5: "mma_done(w0)", 15: "end"}
base = t[0, :, 0][t[0, :, 0] > 0].min()
for k in range(T):
    print(f"--- kernel {k}: clock64 relative to own CTA start | globaltimer ns relative to first traced kernel's first CTA start")
    for slot, n in names.items():
        col = t[k, :, 16 + slot]; ok = col > 0
        if ok.any():
            ck = (t[k, :, 16 + slot] - t[k, :, 16])[ok]
            gts = ""
            if (t[k, :, slot] > 0).any():
                gt = t[k, :, slot][t[k, :, slot] > 0] - base
                gts = f"gt min {gt.min():8.0f} med {np.median(gt):8.0f} max {gt.max():8.0f}"
            print(f"  {n:20s} clk min {ck.min():8.0f} med {np.median(ck):8.0f} max {ck.max():8.0f} | {gts}")
ends = [t[k, :, 15].max() - base for k in range(T)]
print("kernel end-to-end period (ns):", np.diff(ends))
