import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from ternary_vllm_b200 import ops
dev = torch.device("cuda", 0)
flush = torch.empty(int(300e6), dtype=torch.uint8, device=dev)
def timed(fn, it=10):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(it):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort(); return ts[len(ts) // 2]
out = []
for rows in (2144, 2048, 8224, 4096):
    for K in (4096, 1024):
        x = torch.randn(rows, K, device=dev, dtype=torch.bfloat16)
        t = timed(lambda: ops.quant_act_i8(x, kperm=True))
        out.append(f"{rows}x{K}: {t:.1f}us {rows*K*3/t/1e6:.2f}TB/s")
print("WARP_MIN", os.environ.get("TB200_QUANT_WARP_MIN_ROWS"), " | ".join(out), flush=True)
