"""bench_tp.py — the N>1 arm of bench.py (launched under torchrun, one rank per GPU).

Workload = BASELINE.json configs[4]: 7B-class decoder projections 4096x11008, 11008x4096 and the
4096x32000 lm_head, N-sharded across the ranks (ternary_vllm_b200.parallel.shard_rows), decode
batches 1/8/32; every layer = local GEMV on the rank's row block + ONE all-gather of the bf16
output shards over NCCL.  STRONG scaling: the same full layers on more GPUs, so
`value` = algorithmic bytes of the full layers / max-over-ranks device time.
"""
import json
import math
import os
import statistics

TP_SHAPES = [(4096, 11008), (11008, 4096), (4096, 32000)]     # (K, N)
TP_BATCHES = [1, 8, 32]
L2_BYTES = 126e6


def run_tp(args, metric):
    import torch
    import torch.distributed as dist

    from bench import CONFIG_TP, ClockSampler, algorithmic_bytes, peaks
    from ternary_vllm_b200 import ops
    from ternary_vllm_b200.parallel import TPShardedLinear, shard_rows, shard_size

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    pk = peaks()

    # every rank builds only its own row block of every layer
    rings = {}
    for (K, N) in TP_SHAPES:
        a, b = shard_rows(N, world, rank)
        per_layer = shard_size(N, world) * K / 4
        R = max(4, min(32, int(math.ceil(2.2 * L2_BYTES / max(per_layer, 1)))))
        g = torch.Generator(device=dev).manual_seed(1000 + K + N)            # same full matrices on every rank
        layers = []
        for r in range(R):
            # generate the rank's rows only (seeded per layer and row block so that shards are consistent)
            gl = torch.Generator(device=dev).manual_seed(K * 131 + N * 7 + r * 1009 + rank)
            W = torch.randn(b - a, K, generator=gl, device=dev).clip_(-1, 1)
            packed = ops.quantize_pack(W, "ternary")
            scale = torch.full((b - a,), 1.0 / math.sqrt(K), device=dev)
            bias = torch.randn(b - a, generator=gl, device=dev)
            layers.append(TPShardedLinear(packed, scale, bias, K, N, 100.0).enable_peer_gather(32))
        rings[(K, N)] = layers
    cases = [(K, N, M) for (K, N) in TP_SHAPES for M in TP_BATCHES]
    xs = {c: torch.randn(c[2], c[0], device=dev, dtype=torch.bfloat16) for c in cases}

    def step(gather=True):
        """gather=True: the product path (all-gather fused into the GEMV epilogue over NVLink peer memory + one
        cross-rank barrier per layer); gather=False: local compute only; gather="nccl": NCCL all-gather baseline."""
        with torch.no_grad():
            for c in cases:
                for layer in rings[c[:2]]:
                    if gather is True:
                        layer(xs[c])
                    else:
                        y = layer.local_forward(xs[c])
                        if gather == "nccl":
                            layer.gather(y)

    def timed(fn, iters):
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)                                # max over ranks
        dist.barrier()
        return float(t.item())

    for _ in range(max(3, args.warmup)):
        step()
    torch.cuda.synchronize()
    # one step = one CUDA graph (local kernels + NCCL all-gathers): decode layers are a few microseconds,
    # eager launches from Python would measure the host
    n0 = ops.launch_count()
    mode = "cuda graph"
    try:
        if os.environ.get("TB200_TP_GRAPH", "1") != "1":      # the step holds no NCCL call (peer-memory gather), so it captures
            raise RuntimeError("graph capture disabled")
        st = torch.cuda.Stream(dev)
        with torch.cuda.stream(st):
            g_full = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g_full, stream=st):
                step()
            launches_per_step = ops.launch_count() - n0
            g_local = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g_local, stream=st):
                step(False)
        torch.cuda.synchronize()
        run_full, run_local = g_full.replay, g_local.replay
    except Exception as e:                               # capture of the collective refused: eager fallback
        mode = f"eager ({type(e).__name__})"
        torch.cuda.synchronize()
        launches_per_step = None
        run_full, run_local = step, (lambda: step(False))
    for _ in range(3):
        run_full()
    n1 = ops.launch_count()
    sampler = ClockSampler(local)
    sampler.start()
    ms = timed(run_full, args.steps)
    clocks = sampler.stop()
    launches = launches_per_step * args.steps if launches_per_step is not None else ops.launch_count() - n1
    ms_nogather = timed(run_local, max(2, args.steps // 2)) / max(2, args.steps // 2)
    step("nccl")
    ms_nccl = timed(lambda: step("nccl"), 2) / 2            # eager NCCL all-gather baseline (never captured)
    ms_eager = timed(step, 2) / 2                           # the product path launched eagerly, for the same comparison

    # end to end: pinned host activations in, full gathered output back to pinned host memory, per layer
    hx = {c: torch.randn(c[2], c[0], dtype=torch.bfloat16).pin_memory() for c in cases}
    hy = {c: torch.empty(c[2], c[1], dtype=torch.bfloat16).pin_memory() for c in cases}

    def step_e2e():
        with torch.no_grad():
            for c in cases:
                for layer in rings[c[:2]]:
                    xd = hx[c].to(dev, non_blocking=True)
                    y = layer.gather(layer.local_forward(xd))
                    hy[c].copy_(y, non_blocking=True)
        torch.cuda.synchronize()

    step_e2e()
    e2e_steps = max(2, min(5, args.steps))
    ms_e2e = timed(step_e2e, e2e_steps) / e2e_steps
    step_bytes = sum(algorithmic_bytes(*c) * len(rings[c[:2]]) for c in cases)
    h2d = sum(c[2] * c[0] * 2 * len(rings[c[:2]]) for c in cases)
    d2h = sum(c[2] * c[1] * 2 * len(rings[c[:2]]) for c in cases)
    step_tokens = sum(c[2] * len(rings[c[:2]]) for c in cases)
    value = step_bytes * args.steps / (ms * 1e-3) / 1e9
    if rank == 0:
        line = {
            "metric": metric, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "int8 (u8 codes x s8 activations, s32 accumulate)", "data": "synthetic",
            "config": dict(CONFIG_TP, parallelism=f"tp{world} (N-sharded, all-gather)"),
            "tokens_per_s": step_tokens * args.steps / (ms * 1e-3), "clocks": clocks,
            "gpu_launches": launches, "launch_mode": mode, "ms_per_step_without_all_gather": ms_nogather,
            "all_gather_share": max(0.0, 1.0 - ms_nogather / (ms / args.steps)),
            "gather": "fused into the GEMV epilogue: stores to every rank's buffer over NVLink peer memory + 1 barrier per layer",
            "eager_ms_per_step": {"fused_peer_gather": ms_eager, "nccl_all_gather_baseline": ms_nccl},
            "e2e": {"value": step_bytes / (ms_e2e * 1e-3) / 1e9, "unit": "GB/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e,
                    "api": "TPShardedLinear.forward on pinned host activations, gathered output copied back to pinned host"},
            "roofline": {"bound": "hbm", "achieved": value, "peak": pk["hbm_gbs"] * world, "unit": "GB/s",
                                      "frac": value / (pk["hbm_gbs"] * world), "traffic": None,
                                      "note": "aggregate over ranks; decode shards are collective-latency bound"},
            "cpu_baseline": None,
        }
        print(json.dumps(line))
    dist.destroy_process_group()
