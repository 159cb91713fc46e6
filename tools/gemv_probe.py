"""Decode-GEMV timing probe: chain of R distinct layers, M tokens; compares launch modes."""
import math, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ternary_vllm_b200 import ops, _lib as L

dev = torch.device("cuda:0")
K = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
N = int(sys.argv[2]) if len(sys.argv) > 2 else 11008
Ms = [int(a) for a in sys.argv[3].split(",")] if len(sys.argv) > 3 else [1]
R = int(os.environ.get("TB200_PROBE_R", 0)) or max(8, int(math.ceil(2.2 * 126e6 / (N * K / 4))))
g = torch.Generator(device=dev).manual_seed(0)
ld = ops.packed_pitch(K)
packed = torch.empty(R, N, ld, dtype=torch.uint8, device=dev)
for r in range(R):
    packed[r] = ops.quantize_pack(torch.randn(N, K, generator=g, device=dev).clip_(-1, 1), "ternary")
scale = torch.full((R, N), 1 / math.sqrt(K), device=dev)
bias = torch.randn(R, N, generator=g, device=dev)
lib = L.lib()

def timeit(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / reps

for M in Ms:
    x = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
    xs = x.unsqueeze(0).expand(R, M, K).contiguous()
    out = torch.empty(R, M, N, device=dev, dtype=torch.bfloat16)
    bytes_layer = N * K / 4 + M * K * 2 + M * N * 2 + N * 8
    def chain(pdl, fn="matmul"):
        for r in range(R):
            if M <= 8:
                ops.gemv(x, packed[r], scale[r], bias[r], 100.0, pdl=pdl, out=out[r])
            else:
                ops.ternary_matmul(x, packed[r], scale[r], bias[r], 100.0)
    res = {}
    for pdl in (False, True):
        res[f"stream pdl={pdl}"] = timeit(lambda: chain(pdl), 5) / R
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            chain(pdl); s.synchronize()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=s):
                chain(pdl)
        res[f"graph  pdl={pdl}"] = timeit(gr.replay) / R
    if M <= 8:
        def grouped():
            L.check(lib.tb200_gemv_grouped(xs.data_ptr(), L.TB200_BF16, M, K, K, M * K, packed.data_ptr(), ld, N * ld, N,
                                           scale.data_ptr(), bias.data_ptr(), N, 100.0, out.data_ptr(), L.TB200_BF16, N, M * N,
                                           R, 0, torch.cuda.current_stream().cuda_stream), "grouped")
        res["grouped 1 launch"] = timeit(grouped) / R
    ws = torch.empty(lib.tb200_gemv_stream_workspace(M, K, R), dtype=torch.uint8, device=dev)
    try:
        res["stream 2 launches"] = timeit(lambda: ops.gemv_stream(xs, packed, scale, bias, 100.0, out=out, workspace=ws)) / R
        gr2 = torch.cuda.CUDAGraph()
        s2 = torch.cuda.Stream()
        with torch.cuda.stream(s2):
            with torch.cuda.graph(gr2, stream=s2):
                ops.gemv_stream(xs, packed, scale, bias, 100.0, out=out, workspace=ws)
        res["stream graph"] = timeit(gr2.replay) / R
    except Exception as e:
        print("stream:", e)
    for k, t in res.items():
        print(f"K={K} N={N} M={M} R={R} {k:20s} {t*1e6:8.2f} us/layer {bytes_layer/t/1e9:8.1f} GB/s")
