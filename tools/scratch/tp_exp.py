import os, sys, ctypes, math
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torch.distributed as dist
from ternary_vllm_b200 import ops
from ternary_vllm_b200.parallel import TPShardedStream, shard_rows
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device("cuda", local); torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
K, N, R = 4096, 11008, 64
a, b = shard_rows(N, world, rank)
g = torch.Generator(device=dev).manual_seed(1 + rank)
packed = torch.stack([ops.quantize_pack(torch.randn(b - a, K, generator=g, device=dev).clip_(-1, 1), "ternary") for _ in range(R)])
scale = torch.full((R, b - a), 1 / 64.0, device=dev); bias = torch.randn(R, b - a, generator=g, device=dev)
tps = TPShardedStream(packed, scale, bias, K, N, 100.0)
def timed(fn, it=20):
    st = torch.cuda.Stream(dev)
    fn(); torch.cuda.synchronize()
    with torch.cuda.stream(st):
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=st): fn()
    torch.cuda.synchronize()
    for _ in range(3): gr.replay()
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): gr.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it * 1e3
esz = 2
for M in (1, 8, 32):
    xs = torch.randn(R, M, K, device=dev, dtype=torch.bfloat16)
    hdl = tps._hdls[0]
    own = int(hdl.buffer_ptrs[rank]) + rank * tps.n_shard * esz
    oth = int(hdl.buffer_ptrs[(rank + 1) % world]) + rank * tps.n_shard * esz
    loc = torch.empty(R, 32, tps.ld, dtype=torch.bfloat16, device=dev)
    res = {}
    for name, ptrs in (("plain_local", [loc.data_ptr()]), ("own_symm", [own]), ("remote_only", [oth]), ("both", [own, oth])):
        arr = (ctypes.c_void_p * len(ptrs))(*ptrs)
        def fn():
            ops.gemv_stream_tp(xs, packed, scale, bias, 100.0, arr, len(ptrs), torch.bfloat16, tps.ld, 32 * tps.ld, workspace=None)
        ws = torch.empty(ops.L.lib().tb200_gemv_stream_workspace(M, K, R), dtype=torch.uint8, device=dev)
        def fn2():
            ops.gemv_stream_tp(xs, packed, scale, bias, 100.0, arr, len(ptrs), torch.bfloat16, tps.ld, 32 * tps.ld, workspace=ws)
        res[name] = timed(fn2)
    def fb(): tps._hdls[0].barrier(channel=0)
    res["barrier_only"] = timed(fb)
    if rank == 0: print(M, {k: round(v, 1) for k, v in res.items()}, flush=True)
dist.barrier(); dist.destroy_process_group()
