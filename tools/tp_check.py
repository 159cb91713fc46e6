"""Two-rank GPU check of the tensor-parallel layer (run under torchrun on >= 2 GPUs):
NCCL all-gather path and the fused peer-memory gather must both reproduce the single-GPU layer bit for bit.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/tp_check.py
"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import ternary_vllm_b200 as tb
from ternary_vllm_b200.parallel import TPShardedLinear

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
ok = True
for (K, N) in ((4096, 11008), (11008, 4096), (1024, 1000)):
    torch.manual_seed(K + N)                               # the same full layer on every rank
    layer = tb.ScaledTernaryLinear(K, N).to(dev)
    layer.scale.data.fill_(1.0 / K ** 0.5)
    tp = TPShardedLinear.from_layer(layer)
    tp_peer = TPShardedLinear.from_layer(layer).enable_peer_gather(32)
    for M in (1, 8, 32):
        x = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
        dist.broadcast(x, 0)
        with torch.no_grad():
            want = layer(x)
            y1 = tp(x)
            y2 = tp_peer(x).clone()
            y3 = tp_peer(x).clone()                        # second buffer
        torch.cuda.synchronize()
        good = torch.equal(want, y1) and torch.equal(want, y2) and torch.equal(want, y3)
        ok = ok and good
        if rank == 0:
            print(f"K={K} N={N} M={M}: nccl {torch.equal(want, y1)} peer {torch.equal(want, y2)} peer2 {torch.equal(want, y3)}")
    # timing: 50 calls each
    x = torch.randn(1, K, device=dev, dtype=torch.bfloat16)
    for name, mod in (("nccl all-gather", tp), ("fused peer gather", tp_peer)):
        with torch.no_grad():
            for _ in range(5): mod(x)
            torch.cuda.synchronize(); dist.barrier(); t0 = time.perf_counter()
            for _ in range(50): mod(x)
            torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 50
        if rank == 0: print(f"   {name:18s} M=1: {dt*1e6:7.1f} us per layer call (eager, host-inclusive)")
# list form: TPShardedStream (one tb200_gemv_stream_tp call + one barrier for R layers) vs the unsharded tb200_gemv_stream
from ternary_vllm_b200 import ops
from ternary_vllm_b200.parallel import TPShardedStream, shard_rows
for (K, N, R) in ((4096, 11008, 5), (1024, 1000, 3), (11008, 4096, 2)):
    g = torch.Generator(device=dev).manual_seed(K + N + R)      # the same full layers on every rank
    packed = torch.stack([ops.quantize_pack(torch.randn(N, K, generator=g, device=dev).clip_(-1, 1), "ternary") for _ in range(R)])
    scale = torch.rand(R, N, generator=g, device=dev) / K ** 0.5
    bias = torch.randn(R, N, generator=g, device=dev)
    a, b = shard_rows(N, world, rank)
    tps = TPShardedStream(packed[:, a:b].contiguous(), scale[:, a:b].contiguous(), bias[:, a:b].contiguous(), K, N, 100.0)
    for M in (1, 8, 32):
        xs = torch.randn(R, M, K, generator=g, device=dev, dtype=torch.bfloat16)
        want = ops.gemv_stream(xs, packed, scale, bias, 100.0)
        y1 = tps(xs).clone()
        y2 = tps(xs).clone()
        torch.cuda.synchronize()
        good = torch.equal(want, y1) and torch.equal(want, y2)
        ok = ok and good
        if rank == 0:
            print(f"stream K={K} N={N} R={R} M={M}: {good}")
t = torch.tensor([1 if ok else 0], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MIN)
if rank == 0: print("TP CHECK", "PASS" if t.item() == 1 else "FAIL")
dist.destroy_process_group()
sys.exit(0 if t.item() == 1 else 1)
