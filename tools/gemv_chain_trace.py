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
names = {0: "start", 10: "tma_issued", 1: "wait_return", 2: "x_arrived(w0)", 6: "x_arrived(all)", 3: "quantised(w0)", 7: "image_done",
         8: "b_loaded", 4: "tile0_done(w0)", 9: "epi_first_slot", 13: "tile1_done(w0)", 14: "tile2_done(w0)", 5: "consumers_done(w0)", 11: "epi_last_slot_full", 12: "epi_last_summed", 15: "end"}
base = t[0, :, 0][t[0, :, 0] > 0].min()
for k in range(T):
    print(f"--- kernel {k}: globaltimer ns relative to first traced kernel's first CTA start | clock64 relative to own CTA start (median, max)")
    for slot, n in names.items():
        col = t[k, :, 16 + slot]; ok = col > 0
        if ok.any():
            ck = (t[k, :, 16 + slot] - t[k, :, 16])[ok]
            gts = ""
            if (t[k, :, slot] > 0).any():
                gt = t[k, :, slot][t[k, :, slot] > 0] - base
                gts = f"gt min {gt.min():8.0f} med {np.median(gt):8.0f} max {gt.max():8.0f}"
            print(f"  {n:20s} clk min {ck.min():8.0f} med {np.median(ck):8.0f} max {ck.max():8.0f} | {gts}")
    for w, off in (("warp0", 32), ("warp5", 40)):
        c = t[k, :, off:off + 6]
        d = np.diff(c, axis=1)
        print(f"  tile1 {w}: loop_top->half1_issued->fetched->half2_issued->slot_stored->loop_end (median clk):", np.median(d, axis=0), " loop_top rel start:", np.median(c[:, 0] - t[k, :, 16]))
ends = [t[k, :, 15].max() - base for k in range(T)]
print("kernel end-to-end period (ns):", np.diff(ends))
