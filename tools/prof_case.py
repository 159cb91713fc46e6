"""Small fixed workloads for ncu captures (see /opt/skills/guides/B200_PROFILING.md).

    python tools/prof_case.py gemv [M]     # 40 distinct 4096x11008 layers, M tokens (default 1), plain launches
    python tools/prof_case.py gemm         # quant + tcgen05 GEMM, M=4096 K=4096 N=11008, 3 times
    python tools/prof_case.py vit          # the same for the ViT-L QKV shape, M=8224 K=1024 N=3072
"""
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from ternary_vllm_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
case = sys.argv[1] if len(sys.argv) > 1 else "gemv"
g = torch.Generator(device=dev).manual_seed(0)
if case == "gemv":
    M = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    K, N, R = 4096, 11008, 40
    packed = [ops.quantize_pack(torch.randn(N, K, generator=g, device=dev).clip_(-1, 1), "ternary") for _ in range(R)]
    scale = torch.full((N,), 1 / math.sqrt(K), device=dev)
    bias = torch.randn(N, generator=g, device=dev)
    x = torch.randn(M, K, generator=g, device=dev, dtype=torch.bfloat16)
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    for _ in range(2):
        for r in range(R):
            ops.gemv(x, packed[r], scale, bias, 100.0, pdl=True, out=out)
    torch.cuda.synchronize()
elif case == "stream":
    M = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    K, N, R = 4096, 11008, 25
    ld = ops.packed_pitch(K)
    packed = torch.empty(R, N, ld, dtype=torch.uint8, device=dev)
    for r in range(R):
        packed[r] = ops.quantize_pack(torch.randn(N, K, generator=g, device=dev).clip_(-1, 1), "ternary")
    scale = torch.full((R, N), 1 / math.sqrt(K), device=dev)
    bias = torch.randn(R, N, generator=g, device=dev)
    xs = torch.randn(R, M, K, generator=g, device=dev, dtype=torch.bfloat16)
    out = torch.empty(R, M, N, device=dev, dtype=torch.bfloat16)
    for _ in range(3):
        ops.gemv_stream(xs, packed, scale, bias, 100.0, out=out)
    torch.cuda.synchronize()
elif case == "grouped":
    from ternary_vllm_b200 import _lib as L
    M = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    K, N, R = 4096, 11008, 25
    ld = ops.packed_pitch(K)
    packed = torch.empty(R, N, ld, dtype=torch.uint8, device=dev)
    for r in range(R):
        packed[r] = ops.quantize_pack(torch.randn(N, K, generator=g, device=dev).clip_(-1, 1), "ternary")
    scale = torch.full((R, N), 1 / math.sqrt(K), device=dev)
    bias = torch.randn(R, N, generator=g, device=dev)
    xs = torch.randn(R, M, K, generator=g, device=dev, dtype=torch.bfloat16)
    out = torch.empty(R, M, N, device=dev, dtype=torch.bfloat16)
    for _ in range(3):
        L.check(L.lib().tb200_gemv_grouped(xs.data_ptr(), L.TB200_BF16, M, K, K, M * K, packed.data_ptr(), ld, N * ld, N,
                                           scale.data_ptr(), bias.data_ptr(), N, 100.0, out.data_ptr(), L.TB200_BF16, N, M * N,
                                           R, 0, torch.cuda.current_stream().cuda_stream), "grouped")
    torch.cuda.synchronize()
else:
    M, K, N = (32 * 257, 1024, 3072) if case == "vit" else (4096, 4096, 11008)     # vit: configs[3] QKV projection
    packed = ops.quantize_pack(torch.randn(N, K, generator=g, device=dev).clip_(-1, 1), "ternary")
    scale = torch.full((N,), 1 / math.sqrt(K), device=dev)
    bias = torch.randn(N, generator=g, device=dev)
    x = torch.randn(M, K, generator=g, device=dev, dtype=torch.bfloat16)
    for _ in range(3):
        q, s, sq = ops.quant_act_i8(x, kperm=True)
        y = ops.gemm_i8(q, s, sq, K, packed, scale, bias, 100.0)
    torch.cuda.synchronize()
print("ok")
