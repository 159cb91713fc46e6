"""One tb200_chain_forward call on the 4096 -> 11008 -> 4096 stack, for ncu:
    ncu --set full --clock-control none --import-source on -k regex:gemv_chain -s 2 -c 1 -o gpurun_out/chain python tools/chain_profile.py [M]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from chain_bench import build_stack
from ternary_vllm_b200 import ops

M = int(sys.argv[1]) if len(sys.argv) > 1 else 1
dev = torch.device("cuda:0")
d, f, n = 4096, 11008, 16
layers = build_stack(d, f, n, dev, 1)
plan = ops.ChainPlan(layers, d, torch.bfloat16, max_tokens=M)
x = torch.randn(M, d, device=dev, dtype=torch.bfloat16)
for _ in range(4):
    y = plan.forward(x)
torch.cuda.synchronize()
print("ok", float(y.float().abs().mean()))
