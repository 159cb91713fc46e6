"""One launch of every weight-prep / quantiser / 1-bit kernel on the bench shapes, for ncu:
    ncu --set full --clock-control none --import-source on -k regex:"quantize_pack|clip_vec|ternary_pass|ternary_init_row|quant_act_fast|gemv_b1" \
        -c 12 -o gpurun_out/r2_prep python tools/prof_prep.py"""
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from ternary_vllm_b200 import ops

dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(1)
N, K = 11008, 4096
W = torch.randn(N, K, generator=g, device=dev) * 0.02
Wc = W.clip(-1, 1)
flush = torch.empty(int(256e6), dtype=torch.uint8, device=dev)
for fn in (lambda: ops.quantize_pack(Wc, "ternary"), lambda: ops.quantize_pack(Wc.bfloat16(), "ternary"),
           lambda: ops.clip_latent_((W * 60).contiguous()), lambda: ops.ternary_init(W, per_row=True),
           lambda: ops.ternary_init(W, per_row=False)):
    flush.zero_()
    fn()
for (M, Kx) in ((4096, 4096), (8224, 1024)):
    x = torch.randn(M, Kx, generator=g, device=dev, dtype=torch.bfloat16)
    flush.zero_()
    ops.quant_act_i8(x, kperm=True)
p1 = ops.quantize_pack_b1(W)
x = torch.randn(1, K, generator=g, device=dev, dtype=torch.bfloat16)
flush.zero_()
ops.binary_matmul(x, p1, K, torch.full((N,), 1 / math.sqrt(K), device=dev), None, 100.0, False)
torch.cuda.synchronize()
print("ok")
