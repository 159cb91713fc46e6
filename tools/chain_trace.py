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
trace = torch.zeros(148 * 128 * 8, dtype=torch.int64, device=dev)
for _ in range(3):
    plan.forward(x)
torch.cuda.synchronize()
os.environ["TB200_CHAIN_TRACE_PTR"] = str(trace.data_ptr())
plan.forward(x)
torch.cuda.synchronize()
t = trace.cpu().numpy().reshape(148, 128, 8)[:, : 2 * n].astype(np.float64)
t0 = t[t > 0].min()
print("layer  enter(us)  wait_flag  fetch+quant(->image)  mma(team0 last tile)  period | aux: last epilogue done after mma, store+red | last CTA's aux done -> median flag seen   [ns, medians over CTAs]")
for l in range(2 * n):
    v = t[:, l]
    ok = v[:, 0] > 0
    e = v[ok, 0] - t0
    nxt = t[ok, l + 1, 0] - v[ok, 0] if l + 1 < 2 * n else np.zeros(ok.sum())
    aux_after = np.median(v[ok, 4] - v[ok, 3])
    aux_store = np.median(v[ok, 5] - v[ok, 4])
    flag_seen = (np.median(t[ok, l + 1, 1]) - v[ok, 5].max()) if l + 1 < 2 * n else 0
    print(f"  fetch {np.median(v[ok, 6] - v[ok, 1]):5.0f} retries(thread 0) mean {v[ok, 7].mean():4.2f} max {v[ok, 7].max():3.0f}", end="")
    print(f"{l:3d}   {np.median(e) / 1e3:8.2f}   {np.median(v[ok, 1] - v[ok, 0]):7.0f}  {np.median(v[ok, 2] - v[ok, 1]):8.0f}  "
          f"{np.median(v[ok, 3] - v[ok, 2]):10.0f}  {np.median(nxt):10.0f}  |  {aux_after:6.0f} {aux_store:6.0f}  |  {flag_seen:6.0f}   spread(enter) {e.max() - e.min():6.0f}")
