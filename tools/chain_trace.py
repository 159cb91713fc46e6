"""Per-layer timeline of tb200_chain_forward from the experiment library (make ABLATE=1):
    TB200_LIB=ternary_vllm_b200/libternary_b200_ablate.so python tools/chain_trace.py [M]
Prints, per layer, medians over CTAs (ns): wait for the input flag, image build, dot products, and the layer period."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from chain_bench import build_stack
from ternary_vllm_b200 import ops

M = int(sys.argv[1]) if len(sys.argv) > 1 else 1
dev = torch.device("cuda:0")
d, f, n = 4096, 11008, 12
layers = build_stack(d, f, n, dev, 1)
plan = ops.ChainPlan(layers, d, torch.bfloat16, max_tokens=M)
x = torch.randn(M, d, device=dev, dtype=torch.bfloat16)
trace = torch.zeros(148 * 128 * 32, dtype=torch.int64, device=dev)
for _ in range(3):
    plan.forward(x)
torch.cuda.synchronize()
os.environ["TB200_CHAIN_TRACE_PTR"] = str(trace.data_ptr())
plan.forward(x)
torch.cuda.synchronize()
t = trace.cpu().numpy().reshape(148, 128, 32)[:, : 2 * n].astype(np.float64)
t0 = t[t > 0].min()
per = [np.median(t[:, l + 1, 0] - t[:, l, 0]) for l in range(2 * n - 1)]
print("layer period (globaltimer at layer entry, ns, median over CTAs):", " ".join(f"{v:.0f}" for v in per))
print("\nSM-clock deltas (cycles, medians over CTAs): wait | fetch | amax+bar | quantise | bar+load_b | mma phase (of which waiting for tiles, for a slot) || aux of the last row block: wait partials | epilogue math | stores")
for l in range(2 * n):
    v = t[:, l]
    ok = v[:, 8] > 0
    md = lambda a, c: np.median(v[ok, a] - v[ok, c])
    print(f"{l:3d}  {md(9, 8):6.0f} | {md(10, 9):6.0f} | {md(11, 10):6.0f} | {md(14, 11):6.0f} | {md(12, 14):6.0f} | {md(13, 12):6.0f} ({np.median(v[ok, 15]):6.0f}, {np.median(v[ok, 16]):6.0f}) || {md(18, 17):6.0f} | {md(19, 18):6.0f} | {md(20, 19):6.0f}   aux done after mma {md(20, 13):6.0f}")
    rb = " ".join(f"[{np.median(v[ok, a + 1] - v[ok, a]):5.0f} {np.median(v[ok, a + 2] - v[ok, a + 1]):5.0f}]" for a in (1, 4, 21))
    print(f"      row blocks 0-2 [dot products incl. tile wait, slot hand-off]: {rb}   first block starts {np.median(v[ok, 1] - v[ok, 12]):5.0f} after load_b")
    continue
    print(f"{l:3d} "
          f"{md(18, 17):6.0f} | {md(19, 18):6.0f} | {md(20, 19):6.0f}   aux done after mma {md(20, 13):6.0f}")
