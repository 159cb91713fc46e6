"""Worker of tests/test_gpu_tp.py, launched under torchrun with one rank per GPU: every tensor-parallel form must
reproduce the single-GPU layer BIT FOR BIT on real NVLink peer memory.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/tp_worker.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import ternary_vllm_b200 as tb
from ternary_vllm_b200 import ops
from ternary_vllm_b200.parallel import TPShardedLinear, TPShardedStream, shard_rows

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
ok = True


def say(msg):
    if rank == 0:
        print(msg, flush=True)


# per-layer forms: NCCL all-gather (any M, incl. the prefill GEMM) and the fused peer-memory gather (M <= 32);
# N = 16 * (world - 1) leaves the LAST rank with an empty row block (the path choice must stay rank-uniform)
for (K, N) in ((4096, 11008), (11008, 4096), (1024, 1000), (256, 16 * (world - 1) if world > 1 else 16)):
    torch.manual_seed(K + N)                               # the same full layer on every rank
    layer = tb.ScaledTernaryLinear(K, N).to(dev)
    layer.scale.data.fill_(1.0 / K ** 0.5)
    tp = TPShardedLinear.from_layer(layer)
    tp_peer = TPShardedLinear.from_layer(layer).enable_peer_gather(32)
    tp_pre = TPShardedLinear.from_layer(layer).enable_peer_gather(320, multicast="try")     # prefill: NVSwitch multicast stores where available
    tp_pre_u = TPShardedLinear.from_layer(layer).enable_peer_gather(320, multicast=False)   # prefill: one TMA tile store per rank
    for M in (1, 8, 32, 300):
        x = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
        dist.broadcast(x, 0)
        with torch.no_grad():
            want = layer(x)
            y1 = tp(x)
            y2 = tp_peer(x)                                # M <= 32: fused gather (fresh tensor); M = 300: NCCL
            y3 = tp_peer(x, alias_ok=True).clone()         # second buffer, aliased form
            y4 = tp_pre(x)                                 # every M through the fused gather (M = 300: GEMM epilogue)
            y5 = tp_pre_u(x)
        torch.cuda.synchronize()
        good = torch.equal(want, y1) and torch.equal(want, y2) and torch.equal(want, y3) and torch.equal(want, y4) and torch.equal(want, y5)
        ok = ok and good
        say(f"K={K} N={N} M={M}: nccl {torch.equal(want, y1)} peer {torch.equal(want, y2)} peer2 {torch.equal(want, y3)} "
            f"peer-any-M(mc={tp_pre._peer_use_mc}) {torch.equal(want, y4)} peer-any-M(unicast) {torch.equal(want, y5)}")

# list form: ONE tb200_gemv_stream_tp call + one barrier for R layers vs the unsharded tb200_gemv_stream
for (K, N, R) in ((4096, 11008, 5), (1024, 1000, 3), (11008, 4096, 2)):
    g = torch.Generator(device=dev).manual_seed(K + N + R)      # the same full layers on every rank
    packed = torch.stack([ops.quantize_pack(torch.randn(N, K, generator=g, device=dev).clip_(-1, 1), "ternary") for _ in range(R)])
    scale = torch.rand(R, N, generator=g, device=dev) / K ** 0.5
    bias = torch.randn(R, N, generator=g, device=dev)
    a, b = shard_rows(N, world, rank)
    tps = TPShardedStream(packed[:, a:b].contiguous(), scale[:, a:b].contiguous(), bias[:, a:b].contiguous(), K, N, 100.0,
                          multicast="try")          # multicast wherever the box has a mapping
    tps_u = TPShardedStream(packed[:, a:b].contiguous(), scale[:, a:b].contiguous(), bias[:, a:b].contiguous(), K, N, 100.0,
                            multicast=False)
    for M in (1, 8, 32):
        xs = torch.randn(R, M, K, generator=g, device=dev, dtype=torch.bfloat16)
        want = ops.gemv_stream(xs, packed, scale, bias, 100.0)
        y1 = tps(xs)
        y2 = tps(xs, alias_ok=True).clone()
        y3 = tps_u(xs)
        torch.cuda.synchronize()
        good = torch.equal(want, y1) and torch.equal(want, y2) and torch.equal(want, y3)
        ok = ok and good
        say(f"stream K={K} N={N} R={R} M={M}: multicast={tps.multicast} {torch.equal(want, y1) and torch.equal(want, y2)} unicast {torch.equal(want, y3)}")

t = torch.tensor([1 if ok else 0], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MIN)
say("TP CHECK " + ("PASS" if t.item() == 1 else "FAIL") + f" world={world}")
dist.destroy_process_group()
sys.exit(0 if t.item() == 1 else 1)
