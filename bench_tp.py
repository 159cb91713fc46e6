"""bench_tp.py — the N>1 arm of bench.py (launched under torchrun, one rank per GPU).

Headline (`value`): the SAME decode sweep as the N=1 arm (BASELINE.json configs[1]: 4096x4096 and 4096x11008 at
batch 1/8/32), every layer N-sharded across the ranks (ternary_vllm_b200.parallel.shard_rows), over rings of
67 x N / 64 x N distinct layers: the ring grows with the number of ranks so that each rank's shards of it (the same
281 / 721 MB of packed weights as at N=1) stay larger than its L2 and are streamed from HBM.  Per case ONE tb200_gemv_stream_tp call per rank over its row blocks of the
whole ring, whose epilogues store the outputs into every rank's buffer over NVLink peer memory (the all-gather fused
into the kernel), then ONE cross-rank barrier.  WEAK scaling: per-GPU bytes are fixed, the layers are the same full
layers; `value` = algorithmic bytes of the full layers of the step / max-over-ranks device time, directly comparable
with the N=1 line.

`tp_cases` adds configs[4]'s other projections (11008x4096, 4096x32000 lm_head) in the same form; `chain` is the
sequential-layers form (one TPShardedLinear call + one barrier PER LAYER, latency-bound) with its NCCL all-gather
baseline next to it.
"""
import json
import math
import os

TP_EXTRA_SHAPES = [(11008, 4096), (4096, 32000)]     # (K, N): configs[4] beyond the sweep's 4096x11008
CHAIN_SHAPES = [(4096, 11008), (11008, 4096), (4096, 32000)]
CHAIN_LAYERS = 8
L2_BYTES = 126e6


def run_tp(args, metric):
    import torch
    import torch.distributed as dist

    from bench import BATCHES, CONFIG_TP, RING_MIN, SHAPES, ClockSampler, algorithmic_bytes, peaks
    from ternary_vllm_b200 import ops
    from ternary_vllm_b200.parallel import TPShardedLinear, TPShardedStream, shard_rows

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    # stdout carries ONE JSON line: NCCL prints its version banner to file descriptor 1 from C, so everything else written
    # to fd 1 during the run is sent to stderr and the line goes to the saved descriptor at the end
    import sys
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    pk = peaks()

    def build_stream(K, N, R):
        """This rank's row block of R full layers (rows seeded per layer and rank; the reference's own init)."""
        a, b = shard_rows(N, world, rank)
        ld = ops.packed_pitch(K)
        packed = torch.empty((R, b - a, ld), dtype=torch.uint8, device=dev)
        gl = torch.Generator(device=dev).manual_seed(K * 131 + N * 7 + rank)
        for r in range(R):
            packed[r] = ops.quantize_pack(torch.randn(b - a, K, generator=gl, device=dev).clip_(-1, 1), "ternary")
        scale = torch.full((R, b - a), 1.0 / math.sqrt(K), device=dev)
        bias = torch.randn((R, b - a), generator=gl, device=dev)
        return TPShardedStream(packed, scale, bias, K, N, 100.0, max_tokens=32)

    streams, ring_len = {}, {}
    for (K, N) in SHAPES + TP_EXTRA_SHAPES:
        # the N=1 ring length x world: every rank's SHARDS of the ring stay larger than its own L2 (the timing rule) and
        # the packed bytes a GPU streams per step are the same at every N -> weak scaling
        R = world * max(RING_MIN if (K, N) in SHAPES else 24, int(math.ceil(2.2 * L2_BYTES / (N * K / 4))))
        ring_len[(K, N)] = R
        streams[(K, N)] = build_stream(K, N, R)
    cases = [(K, N, M) for (K, N) in SHAPES for M in BATCHES]
    extra = [(K, N, M) for (K, N) in TP_EXTRA_SHAPES for M in BATCHES]
    g = torch.Generator(device=dev).manual_seed(99)                 # same activations on every rank (x is replicated)
    xs = {c: torch.randn(ring_len[c[:2]], c[2], c[0], generator=g, device=dev, dtype=torch.bfloat16) for c in cases + extra}

    def run_case(c, barrier=True):
        return streams[c[:2]](xs[c], barrier=barrier, alias_ok=True)

    def step(barrier=True):
        for c in cases:
            run_case(c, barrier)

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

    def graphed(fn):
        """fn captured as one CUDA graph (the step holds no NCCL call: peer stores + symmetric-memory barrier)."""
        st = torch.cuda.Stream(dev)
        fn()
        torch.cuda.synchronize()
        n0 = ops.launch_count()
        with torch.cuda.stream(st):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=st):
                fn()
        torch.cuda.synchronize()
        return gr, ops.launch_count() - n0

    for _ in range(max(3, args.warmup)):
        step()
    torch.cuda.synchronize()
    mode = "cuda graph"
    try:
        if os.environ.get("TB200_TP_GRAPH", "1") != "1":
            raise RuntimeError("graph capture disabled")
        g_full, launches_per_step = graphed(step)
        g_local, _ = graphed(lambda: step(False))
        run_full, run_local = g_full.replay, g_local.replay
    except Exception as e:                               # capture refused: eager fallback
        mode = f"eager ({type(e).__name__}: {e})"[:200]
        torch.cuda.synchronize()
        n0 = ops.launch_count()
        step()
        launches_per_step = ops.launch_count() - n0
        run_full, run_local = step, (lambda: step(False))
    for _ in range(3):
        run_full()
    sampler = ClockSampler(local)
    sampler.start()
    ms = timed(run_full, args.steps)
    clocks = sampler.stop()
    n_loc = max(2, args.steps // 2)
    ms_nobarrier = timed(run_local, n_loc) / n_loc

    # per case (sweep + configs[4] extras): one call + one barrier, replayed as a graph
    def case_rows(cs):
        rows = []
        for c in cs:
            K, N, M = c
            R = ring_len[(K, N)]
            try:
                gr, _ = graphed(lambda: run_case(c))
                fn = gr.replay
            except Exception:
                fn = lambda: run_case(c)        # noqa: E731
            for _ in range(2):
                fn()
            t = timed(fn, 10) / 10 * 1e-3
            b = algorithmic_bytes(K, N, M) * R
            rows.append({"K": K, "N": N, "M": M, "layers": R, "us_per_call": t * 1e6, "us_per_layer": t / R * 1e6,
                         "gbs": b / t / 1e9, "frac_hbm_measured_aggregate": b / t / 1e9 / (pk["hbm_gbs"] * world),
                         "tokens_per_s": M * R / t, "gather_bytes_received_per_rank": R * M * N * 2 * (world - 1) // world})
        return rows

    per_case = case_rows(cases)
    tp_cases = case_rows(extra)

    # the sequential-layers form: one call + one barrier per layer (latency-bound), NCCL all-gather baseline beside it
    chain = []
    xs1 = {}
    for (K, N) in CHAIN_SHAPES:
        st = streams[(K, N)]
        layers = [TPShardedLinear(st.packed[r], st.scale[r], st.bias[r], K, N, 100.0).enable_peer_gather(32)
                  for r in range(CHAIN_LAYERS)]
        for M in BATCHES:
            x = xs1.setdefault((K, M), torch.randn(M, K, device=dev, dtype=torch.bfloat16))

            def run_chain(mode_="peer"):
                with torch.no_grad():
                    for layer in layers:
                        if mode_ == "peer":
                            layer(x, alias_ok=True)
                        else:
                            layer.gather(layer.local_forward(x))
            try:
                gr, _ = graphed(run_chain)
                fn = gr.replay
            except Exception:
                fn = run_chain
            fn()
            t_peer = timed(fn, 5) / 5 / CHAIN_LAYERS * 1e-3
            run_chain("nccl")
            t_nccl = timed(lambda: run_chain("nccl"), 3) / 3 / CHAIN_LAYERS * 1e-3
            b = algorithmic_bytes(K, N, M)
            chain.append({"K": K, "N": N, "M": M, "us_per_layer_fused_peer_gather": t_peer * 1e6, "gbs": b / t_peer / 1e9,
                          "us_per_layer_nccl_all_gather_eager": t_nccl * 1e6})
        del layers

    # ---- strong scaling: the FIXED N=1 rings (64 / 67 layers), every layer N-sharded: total work does not grow with the ranks.
    # A rank's shards of a ring fall below its L2 at 4 / 8 ranks, so L2 is flushed (256 MB memset) before every timed call.
    flush = torch.empty(int(256e6), dtype=torch.uint8, device=dev)
    strong = []
    for c in cases:
        K, N, M = c
        R0 = ring_len[(K, N)] // world
        st0 = streams[(K, N)]
        sub = TPShardedStream(st0.packed[:R0], st0.scale[:R0], st0.bias[:R0], K, N, 100.0, max_tokens=32)
        x0 = xs[c][:R0]
        for _ in range(2):
            sub(x0, alias_ok=True)
        ts = []
        for _ in range(5):
            flush.zero_()
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            sub(x0, alias_ok=True)
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ts.append(float(t.item()))
        t = sorted(ts)[len(ts) // 2] * 1e-3
        b = algorithmic_bytes(K, N, M) * R0
        strong.append({"K": K, "N": N, "M": M, "layers": R0, "us_per_call": t * 1e6, "gbs": b / t / 1e9, "tokens_per_s": M * R0 / t,
                       "l2_policy": "L2 flushed (256 MB memset) before every timed call; eager launch + barrier"})
        del sub

    # ---- prefill (configs[4]: M = 4096 tokens, N-sharded): quantiser + tcgen05 GEMM per shard, then the gather --
    # (a) fused: the GEMM epilogue's TMA tile stores write the shard into every rank's buffer over NVLink + one barrier;
    # (b) baseline: local GEMM, then NCCL all_gather_into_tensor + the re-layout of the rank-major result to [M, N]
    prefill = []
    Mp = 4096
    xp = {}
    for (K, N) in CHAIN_SHAPES:
        st0 = streams[(K, N)]
        x = xp.setdefault(K, torch.randn(Mp, K, device=dev, dtype=torch.bfloat16))
        dist.broadcast(x, 0)
        lay_f = TPShardedLinear(st0.packed[0], st0.scale[0], st0.bias[0], K, N, 100.0).enable_peer_gather(Mp, multicast="try")
        lay_n = TPShardedLinear(st0.packed[0], st0.scale[0], st0.bias[0], K, N, 100.0)
        lay_u = TPShardedLinear(st0.packed[0], st0.scale[0], st0.bias[0], K, N, 100.0).enable_peer_gather(Mp, multicast=False)

        def t_of(fn, reps=5):
            fn()
            ts = []
            for _ in range(reps):
                flush.zero_()
                dist.barrier()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                fn()
                e1.record()
                torch.cuda.synchronize()
                tt = torch.tensor([e0.elapsed_time(e1)], device=dev)
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                ts.append(float(tt.item()))
            return sorted(ts)[len(ts) // 2] * 1e-3
        with torch.no_grad():
            try:                                     # the fused call holds no NCCL collective: replayed as a CUDA graph
                gr_f, _ = graphed(lambda: lay_f(x, alias_ok=True))
                t_fused = t_of(gr_f.replay)
                fused_mode = "cuda graph"
            except Exception as e:
                t_fused = t_of(lambda: lay_f(x, alias_ok=True))
                fused_mode = f"eager ({type(e).__name__})"
            try:
                gr_u, _ = graphed(lambda: lay_u(x, alias_ok=True))
                t_unicast = t_of(gr_u.replay)
            except Exception:
                t_unicast = t_of(lambda: lay_u(x, alias_ok=True))
            n_loc = lay_n.stop - lay_n.start
            y_loc = torch.empty(Mp, max(n_loc, 1), device=dev, dtype=torch.bfloat16)
            gr_l, _ = graphed(lambda: ops.ternary_matmul(x, lay_n.packed, lay_n.scale, lay_n.bias, 100.0, out=y_loc[:, :n_loc]) if n_loc else None)
            t_local = t_of(gr_l.replay)
            yl = lay_n.local_forward(x)
            t_gather = t_of(lambda: lay_n.gather(yl))
            t_nccl = t_of(lambda: lay_n(x))
        ops_ = 2.0 * Mp * N * K
        prefill.append({"K": K, "N": N, "M": Mp, "us_fused_peer_gather": t_fused * 1e6, "tops_fused": ops_ / t_fused / 1e12,
                        "fused_launch_mode": fused_mode, "fused_multicast": bool(lay_f._peer_use_mc),
                        "default_form": "multicast" if (world > 4 and lay_f._peer_use_mc) else "unicast",
                        "us_fused_default": (t_fused if (world > 4 and lay_f._peer_use_mc) else t_unicast) * 1e6,
                        "us_fused_unicast_tma_stores": t_unicast * 1e6, "us_local_gemm_only": t_local * 1e6, "us_nccl_all_gather_plus_relayout": t_gather * 1e6,
                        "us_local_gemm_then_nccl": t_nccl * 1e6, "tops_nccl": ops_ / t_nccl / 1e12,
                        "tokens_per_s_fused": Mp / t_fused,
                        "gather_bytes_received_per_rank": Mp * N * 2 * (world - 1) // world})
        del lay_f, lay_n, lay_u
        torch.cuda.empty_cache()

    # end to end: pinned host activations in, full gathered outputs back to pinned host memory, per case
    hx = {c: torch.randn(ring_len[c[:2]], c[2], c[0], dtype=torch.bfloat16).pin_memory() for c in cases}
    hy = {c: torch.empty(ring_len[c[:2]], c[2], c[1], dtype=torch.bfloat16).pin_memory() for c in cases} if rank == 0 else {}
    dx = {c: torch.empty_like(xs[c]) for c in cases}

    def step_e2e():
        for c in cases:
            dx[c].copy_(hx[c], non_blocking=True)
            y = streams[c[:2]](dx[c], alias_ok=True)
            if rank == 0:                       # the caller reads the gathered result once, on rank 0
                hy[c].copy_(y, non_blocking=True)
        torch.cuda.synchronize()

    step_e2e()
    e2e_steps = max(3, min(10, args.steps))
    ms_e2e = timed(step_e2e, e2e_steps) / e2e_steps
    step_bytes = sum(algorithmic_bytes(*c) * ring_len[c[:2]] for c in cases)
    h2d = sum(hx[c].numel() * 2 for c in cases)
    d2h = sum(ring_len[c[:2]] * c[2] * c[1] * 2 for c in cases)
    step_tokens = sum(c[2] * ring_len[c[:2]] for c in cases)
    value = step_bytes * args.steps / (ms * 1e-3) / 1e9
    if rank == 0:
        line = {
            "metric": metric, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int8 (u8 codes x s8 activations, s32 accumulate)", "data": "synthetic",
            "config": dict(CONFIG_TP, parallelism=f"tp{world} (N-sharded, all-gather fused into the GEMV epilogues)"),
            "tokens_per_s": step_tokens * args.steps / (ms * 1e-3), "clocks": clocks,
            "gpu_launches": launches_per_step * args.steps, "launches_per_step": launches_per_step, "launch_mode": mode,
            "ms_per_step_without_barriers": ms_nobarrier, "multicast_stores": bool(streams[SHAPES[0]].multicast),
            "barrier_share": max(0.0, 1.0 - ms_nobarrier / (ms / args.steps)),
            "gather": "fused into the epilogues of tb200_gemv_stream_tp: stores to every rank's buffer over NVLink peer "
                      "memory + ONE symmetric-memory barrier per call (a call = one sweep case over its ring of layers)",
            "e2e": {"value": step_bytes / (ms_e2e * 1e-3) / 1e9, "unit": "GB/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e,
                    "api": "TPShardedStream.forward per sweep case on pinned host activations (every rank copies x in; "
                           "rank 0 copies the full gathered output back to pinned host memory)"},
            "cases": per_case, "tp_cases": tp_cases, "chain": chain, "strong_scaling": strong, "prefill_tp": prefill,
            "roofline": {"bound": "hbm", "achieved": value, "peak": pk["hbm_gbs"] * world, "unit": "GB/s",
                         "frac": value / (pk["hbm_gbs"] * world), "traffic": None, "peak_source": pk["source"],
                         "note": "aggregate over ranks (peak = ranks x measured copy rate); per-rank kernel as in the N=1 line"},
            "cpu_baseline": None,
        }
        os.write(real_stdout, (json.dumps(line) + "\n").encode())
    dist.barrier()
    dist.destroy_process_group()
