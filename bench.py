#!/usr/bin/env python
"""bench.py — throughput of the ternary-Linear hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

N=1 workload = BASELINE.json configs[1], the decode GEMV sweep: ternary Linear 4096x4096 and
4096x11008 at batch 1/8/32 with bf16 activations.  One STEP runs every case of the sweep over a
ring of >= 64 distinct layers whose packed weights exceed L2 (so every layer is streamed from HBM),
one tb200_gemv_stream call per case (activation quantiser launch + ONE persistent GEMV launch for
the whole ring; tcgen05 grouped kernel above 8 tokens).  `value` = algorithmic bytes of the step /
device time (GB/s), the step replayed as one CUDA graph whose six independent cases sit on three branches
(`step_schedule`; the one-stream figure is reported beside it); `e2e` = the same sweep through the
host-buffer C-ABI call (tb200_gemv_stream_host: H2D of all activations of the ring, the two
launches, D2H of all outputs, per case).  Every case also reports the same ring as a CHAIN of
per-layer calls (`cases[i].chain`: tb200_gemv with programmatic dependent launch / tb200_linear_forward),
which is what the sequential layers of one model do and is latency-bound.
Extra keys: the tcgen05 prefill GEMM (configs[2]) in TOPS, ViT-L shapes (configs[3]), `roofline`,
`cpu_baseline` (the oracle port of the reference forward timed on this box's host cores).

N>1 (torchrun -> bench_tp.py): the same sweep with every layer N-sharded across the ranks (configs[4]'s
tensor parallelism), the all-gather fused into the GEMV epilogues over NVLink peer memory + one barrier per
call (weak scaling: the ring of full layers grows with the ranks so that every rank's shards stay larger than
its L2); configs[4]'s other projections in `tp_cases`.
"""
import argparse
import json
import math
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
    os.environ["NCCL_DEBUG"] = "WARN"        # NCCL prints its version banner on stdout; stdout carries one JSON line

SHAPES = [(4096, 4096), (4096, 11008)]          # (K, N)
BATCHES = [1, 8, 32]
L2_BYTES = 126e6


def algorithmic_bytes(K, N, M):
    """packed weights + bf16 activations + bf16 outputs + scale/bias."""
    return N * K / 4 + M * K * 2 + M * N * 2 + N * 8


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm_gbs": p["hbm_gbs"], "bf16_tflops": p["bf16_tflops"], "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "source": "fallback (B200_PROFILING.md)"}


# ------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40,
                 "sw_thermal_slowdown": 0x20, "hw_power_brake": 0x80}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        if self.nv:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()

    def stop(self):
        if self._thr:
            self._stop.set()
            self._thr.join()
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------------------------------------ reference arm
_CPU_DATA = {}


def _cpu_case(K, N, M):
    """Synthetic layer for the CPU arm, built once (outside every timed region)."""
    import torch
    key = (K, N, M)
    if key not in _CPU_DATA:
        g = torch.Generator().manual_seed(1234)
        W = torch.randn(N, K, generator=g).clip(-1, 1)
        x = torch.randn(M, K, generator=g)
        _CPU_DATA[key] = (W, x, torch.full((N,), 1.0 / K ** 0.5), torch.randn(N, generator=g))
    return _CPU_DATA[key]


def cpu_forward_sample(cases, iters, warm, kind="c1"):
    """Times the reference forward on the host cores, one layer per case; per-case time = MEDIAN of `iters` runs after
    `warm` warm-ups.  kind "c1": the oracle's restatement of ScaledTernaryLinear.forward as shipped
    (binary/modules.py:101-109, INCLUDING the per-call re-ternarisation of the latent weights); kind "c2": the same
    forward with the weights ternarised once outside the timed region (F.linear(x, q) * scale + bias, clip) -- the fair
    matmul-only baseline.  Returns (bytes, seconds, per-case list)."""
    import torch
    from oracle import ref_modules as om
    torch.set_num_threads(os.cpu_count() or 1)
    total_b, total_t, per = 0.0, 0.0, []
    for (K, N, M) in cases:
        W, x, scale, bias = _cpu_case(K, N, M)
        if kind == "c2":
            q = om.ternary_codes(W).to(torch.float32)
            fn = lambda: om.packed_ternary_matmul(x, q, scale, bias, 100.0)
        else:
            fn = lambda: om.scaled_linear_forward(x.clone(), W, scale, bias, 100.0, "ternary")
        for _ in range(warm):
            fn()
        ts = []
        for _ in range(iters):
            t0 = time.perf_counter()
            fn()
            ts.append(time.perf_counter() - t0)
        dt = statistics.median(ts)
        b = algorithmic_bytes(K, N, M)
        per.append({"K": K, "N": N, "M": M, "ms": dt * 1e3, "gbs": b / dt / 1e9, "tokens_per_s": M / dt})
        total_b += b
        total_t += dt
    return total_b, total_t, per


def cpu_other_baselines(iters=5, warm=2):
    """the reference's own micro-benchmark `ternary_mm` at 1024^2 next to torch.matmul
    (matrix_mul_tests.py:6-41) and the weight-prep baseline linear_to_quantized(..., "ternary") on 4096 x 11008
    (binary/modules.py:275-298), oracle ports on the host cores, median of `iters` after `warm` warm-ups."""
    import torch
    from oracle import ref_modules as om
    torch.set_num_threads(os.cpu_count() or 1)

    def med(fn, n_it=iters, n_warm=warm):
        for _ in range(n_warm):
            fn()
        ts = []
        for _ in range(n_it):
            t0 = time.perf_counter()
            fn()
            ts.append(time.perf_counter() - t0)
        return statistics.median(ts)
    g = torch.Generator().manual_seed(5)
    m = torch.randn(1024, 1024, generator=g)
    tm = torch.randint(-1, 2, (1024, 1024), generator=g).float()
    t_tmm = med(lambda: om.ternary_mm(m, tm))
    t_mm = med(lambda: torch.matmul(m, tm.T))
    W = torch.randn(11008, 4096, generator=g) * 0.02
    t_init = med(lambda: om.ternary_init(W, "uniform"), n_it=max(3, iters - 2), n_warm=1)
    return {"c3_ternary_mm_1024": {"ms": t_tmm * 1e3, "dense_matmul_ms": t_mm * 1e3,
                                   "what": "ternary_mm (matrix_mul_tests.py:29-37) vs torch.matmul, 1024x1024 fp32"},
            "c4_linear_to_quantized_ternary_4096x11008": {"ms": t_init * 1e3, "gbs_one_pass": 4096 * 11008 * 4 / t_init / 1e9,
                                                          "what": "ternary branch of linear_to_quantized (binary/modules.py:275-298): "
                                                                  "10-iteration thresholded mean + latent rescale, uniform scale, fp32"}}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    cases = [(K, N, M) for (K, N) in SHAPES for M in BATCHES]
    for c in cases:
        _cpu_case(*c)
    # every step = one forward of one layer per case (bounded sample of the sweep)
    for _ in range(args.warmup):
        cpu_forward_sample(cases, 1, 0)
    t0 = time.perf_counter()
    tot_b = 0.0
    per = None
    for _ in range(args.steps):
        b, _, per = cpu_forward_sample(cases, 1, 0)
        tot_b += b
    dt = time.perf_counter() - t0
    val = tot_b / dt / 1e9
    cores = torch.get_num_threads()
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "GB/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": CONFIG_N1 if args.gpus == 1 else CONFIG_TP,
        "cpu_baseline": {"value": val, "unit": "GB/s", "cores": cores, "kind": "port",
                         "sample": "one layer per sweep case per step (6 forwards of configs[1]), oracle port of "
                                   "ScaledTernaryLinear.forward incl. per-call re-ternarisation, torch CPU fp32"},
        "e2e": {"value": val, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "cases": per, "gpu_launches": 0,
    }
    print(json.dumps(line))


METRIC = "packed-ternary decode GEMV sweep, algorithmic HBM bytes / time"
CONFIG_N1 = {"workload": "configs[1] decode GEMV sweep: ternary Linear 4096x4096 and 4096x11008, batch 1/8/32, "
                         "bf16 activations, ring of distinct layers per case",
             "l2_policy": "inputs larger than L2 (ring of 67 x 4.2 MB / 64 x 11.3 MB distinct packed layers per case, "
                          "every layer read once per step)",
             "api": "tb200_gemv_stream: one call per sweep case over its ring of layers"}
CONFIG_TP = {"workload": "configs[1] decode GEMV sweep (4096x4096, 4096x11008, batch 1/8/32, the N=1 rings) with every layer "
                         "N-sharded across the ranks as in configs[4]; configs[4]'s 11008x4096 and 4096x32000 in tp_cases",
             "l2_policy": "inputs larger than L2 on every rank (rings of 67 x ranks / 64 x ranks distinct full layers: each rank "
                          "streams the same 281 / 721 MB of packed shards per case as the single GPU does, every layer read once per step)",
             "api": "tb200_gemv_stream_tp: one call per sweep case over the rank's row blocks of the ring + one barrier"}


# ------------------------------------------------------------------------------------ B200 arm
class Ring:
    """R distinct ternary layers of one shape, stored contiguously (so grouped launches work too)."""

    def __init__(self, K, N, R, dev, seed):
        import torch
        from ternary_vllm_b200 import ops
        self.K, self.N, self.R = K, N, R
        ld = ops.packed_pitch(K)
        self.packed = torch.empty((R, N, ld), dtype=torch.uint8, device=dev)
        g = torch.Generator(device=dev).manual_seed(seed)
        for r in range(R):
            W = torch.randn(N, K, generator=g, device=dev).clip_(-1, 1)     # the reference's own init (modules.py:96)
            self.packed[r] = ops.quantize_pack(W, "ternary")
        self.scale = torch.full((R, N), 1.0 / math.sqrt(K), device=dev)
        self.bias = torch.randn((R, N), generator=g, device=dev)


def run_b200(args):
    import torch
    import ternary_vllm_b200 as tb
    from ternary_vllm_b200 import ops, _lib as L

    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        from bench_tp import run_tp
        return run_tp(args, METRIC)
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    lib = L.lib()
    pk = peaks()

    rings = {}
    for (K, N) in SHAPES:
        R = max(RING_MIN, int(math.ceil(2.2 * L2_BYTES / (N * K / 4))))
        rings[(K, N)] = Ring(K, N, R, dev, seed=K + N)
    cases = [(K, N, M) for (K, N) in SHAPES for M in BATCHES]
    g = torch.Generator(device=dev).manual_seed(99)
    xs = {c: torch.randn(rings[c[:2]].R, c[2], c[0], generator=g, device=dev, dtype=torch.bfloat16) for c in cases}
    outs = {c: torch.empty(rings[c[:2]].R, c[2], c[1], device=dev, dtype=torch.bfloat16) for c in cases}
    wss = {c: torch.empty(lib.tb200_gemv_stream_workspace(c[2], c[0], rings[c[:2]].R), dtype=torch.uint8, device=dev) for c in cases}

    def launch_case(c):
        """One decode pass of case c over its ring of layers: tb200_gemv_stream (activation quantiser launch + ONE
        persistent GEMV launch for the whole ring)."""
        ring = rings[c[:2]]
        ops.gemv_stream(xs[c], ring.packed, ring.scale, ring.bias, 100.0, out=outs[c], workspace=wss[c])

    def launch_chain(c):
        """The same ring as a chain of per-layer calls (what a model's sequential layers do): tb200_gemv with
        programmatic dependent launch up to 8 tokens, tb200_linear_forward above."""
        K, N, M = c
        ring = rings[(K, N)]
        for r in range(ring.R):
            if M <= 8:
                ops.gemv(xs[c][r], ring.packed[r], ring.scale[r], ring.bias[r], 100.0, pdl=True, out=outs[c][r])
            else:
                ops.ternary_matmul(xs[c][r], ring.packed[r], ring.scale[r], ring.bias[r], 100.0, out=outs[c][r])

    def launch_step():
        for c in cases:
            launch_case(c)

    # The six cases are independent problems: captured on three branches of the graph, one case's quantiser launch and the ramp of
    # its persistent kernel run in the tail of the previous case's kernel (every kernel here is one CTA per SM, so the
    # branches take turns on the SMs rather than share them).  TB200_BENCH_FORK=0 captures them back to back on one stream.
    fork = int(os.environ.get("TB200_BENCH_FORK", "3"))      # measured: 1 branch 4.78, 2: 4.85, 3: 4.96, 4: 4.97, 6: 4.94 TB/s
    side_streams = [torch.cuda.Stream(dev) for _ in range(max(0, fork - 1))]

    def launch_step_forked():
        main = torch.cuda.current_stream(dev)
        branches = [main] + side_streams
        ev = torch.cuda.Event()
        ev.record(main)
        for b in side_streams:
            b.wait_event(ev)
        # longest cases first on alternating branches
        order = sorted(cases, key=lambda c: -algorithmic_bytes(*c) * rings[c[:2]].R)
        for i, c in enumerate(order):
            with torch.cuda.stream(branches[i % len(branches)]):
                launch_case(c)
        for b in side_streams:
            e = torch.cuda.Event()
            e.record(b)
            main.wait_event(e)

    step_bytes = sum(algorithmic_bytes(*c) * rings[c[:2]].R for c in cases)
    step_tokens = sum(c[2] * rings[c[:2]].R for c in cases)

    def timed_graph(fn, stream, reps, warm=3):
        fn()
        stream.synchronize()
        n0 = ops.launch_count()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=stream):
            fn()
        launches = ops.launch_count() - n0
        for _ in range(warm):
            gr.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        stream.synchronize()
        e0.record(stream)
        for _ in range(reps):
            gr.replay()
        e1.record(stream)
        stream.synchronize()
        return e0.elapsed_time(e1) * 1e-3 / reps, launches, gr

    # ---- one step = all six cases, captured as one CUDA graph ----
    stream = torch.cuda.Stream(dev)
    step_fn = launch_step_forked if fork > 1 else launch_step
    with torch.cuda.stream(stream):
        launch_step()
        stream.synchronize()
        t_serial, _, _ = timed_graph(launch_step, stream, max(3, min(args.steps, 20)))      # reported beside the forked step
        n0 = ops.launch_count()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            step_fn()
        launches_per_step = ops.launch_count() - n0
        for _ in range(max(3, args.warmup)):
            graph.replay()
        stream.synchronize()
        sampler = ClockSampler(local)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sampler.start()
        torch.cuda.synchronize()
        ev0.record(stream)
        for _ in range(args.steps):
            graph.replay()
        ev1.record(stream)
        torch.cuda.synchronize()
        clocks = sampler.stop()
        ms = ev0.elapsed_time(ev1)
    value = step_bytes * args.steps / (ms * 1e-3) / 1e9

    # ---- per-case device numbers: the streaming call, and the per-layer chain next to it ----
    per_case = []
    with torch.cuda.stream(stream):
        for c in cases:
            K, N, M = c
            R = rings[(K, N)].R
            t_call, nl, _ = timed_graph(lambda: launch_case(c), stream, 20)
            t_chain, nlc, _ = timed_graph(lambda: launch_chain(c), stream, 5)
            t = t_call / R
            b = algorithmic_bytes(K, N, M)
            per_case.append({"K": K, "N": N, "M": M, "layers": R, "us_per_layer": t * 1e6, "us_per_call": t_call * 1e6,
                             "launches_per_call": nl, "gbs": b / t / 1e9, "frac_hbm_measured": b / t / 1e9 / pk["hbm_gbs"],
                             "frac_hbm_spec_8TBs": b / t / 1e9 / 8000.0,
                             "tokens_per_s": M / t,
                             "chain": {"us_per_layer": t_chain / R * 1e6, "launches_per_layer": nlc / R,
                                       "gbs": b / (t_chain / R) / 1e9,
                                       "api": "one call per layer (tb200_gemv + PDL up to 8 tokens, tb200_linear_forward above)"}})
    dom = [p_ for p_ in per_case if p_["N"] == 11008 and p_["M"] == 1][0]
    roofline = {"bound": "hbm",
                "kernel": "gemv_stream_kernel<NG=1,BREG,TEAMS=2> (+ quant_image_kernel) on %d layers of 4096x11008, M=1" % dom["layers"],
                "achieved": dom["gbs"], "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": dom["gbs"] / pk["hbm_gbs"],
                "peak_source": pk["source"], "frac_of_spec_8TBs": dom["gbs"] / 8000.0,
                "algorithmic_bytes_per_launch": algorithmic_bytes(4096, 11008, 1) * dom["layers"],
                "avg_launch_us": dom["us_per_call"], "traffic": TRAFFIC_BYTES_PER_LAYER * dom["layers"],
                "chain_of_single_layer_launches_gbs": dom["chain"]["gbs"]}

    # ---- e2e: host buffers through the C ABI ----
    e2e = run_e2e(args, lib, dev, rings, cases, step_bytes)

    # ---- tcgen05 prefill GEMM (configs[2]) ----
    gemm = run_gemm(dev, pk) if not args.no_gemm else None

    # ---- CPU baseline: bounded sample on this box's host cores ----
    cb = None
    if not args.no_cpu:
        b, t, per = cpu_forward_sample(cases, 5, 2, "c1")
        b2, t2, per2 = cpu_forward_sample(cases, 5, 2, "c2")
        cb = {"value": b / t / 1e9, "unit": "GB/s", "cores": torch.get_num_threads(), "kind": "port",
              "sample": "C1: one layer per sweep case, median of 5 timed forwards after 2 warm-ups each (oracle port of the "
                        "reference forward incl. its per-call re-ternarisation, torch CPU fp32)",
              "cases": per,
              "c2_pre_ternarised": {"value": b2 / t2 / 1e9, "unit": "GB/s", "cases": per2,
                                    "what": "F.linear(x, q) * scale + bias, clip with q ternarised once (matmul-only baseline), "
                                            "same cases, median of 5 after 2 warm-ups"},
              "gpu_over_c1": {"device": value / (b / t / 1e9), "e2e": e2e["value"] / (b / t / 1e9)},
              "gpu_over_c2": {"device": value / (b2 / t2 / 1e9), "e2e": e2e["value"] / (b2 / t2 / 1e9)}}
        cb.update(cpu_other_baselines())

    line = {
        "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": 1, "steps": args.steps, "warmup": max(3, args.warmup),
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int8 (u8 codes x s8 activations, s32 accumulate)", "data": "synthetic",
        "config": CONFIG_N1, "tokens_per_s": step_tokens * args.steps / (ms * 1e-3),
        "frac_hbm_measured": value / pk["hbm_gbs"], "clocks": clocks,
        "step_schedule": {"branches": max(1, fork), "note": "the six independent cases captured on this many branches of one CUDA graph "
                          "(longest first, alternating): a case's quantiser and the ramp of its persistent kernel overlap the tail of "
                          "the previous case's kernel", "value_one_stream": step_bytes / t_serial / 1e9, "ms_per_step_one_stream": t_serial * 1e3},
        "e2e": e2e, "gpu_launches": launches_per_step * args.steps, "launches_per_step": launches_per_step,
        "cases": per_case, "roofline": roofline, "mlp_chain": None if args.no_extra else run_mlp(dev, pk),
        "prep": None if args.no_extra else run_prep(dev, pk), "binary_1bit": None if args.no_extra else run_binary_1bit(dev, pk), "gemm": gemm, "vit_l14": None if args.no_gemm else run_vit(dev),
        "cpu_baseline": cb,
    }
    print(json.dumps(line))


RING_MIN = 64               # layers per ring (>= 2.2 x L2 in packed bytes for both shapes)
# dram__bytes_read.sum + dram__bytes_write.sum of gemv_stream_kernel<1,1,2,2,0> per layer of 4096x11008, M=1:
# (284.23 + 3.86) MB over the 25 layers of the ncu capture (profiles/r2_gemv_stream_m1.summary.txt, the kernel with the
# contiguous block assignment); the algorithmic figure is 11.39 MB per layer
TRAFFIC_BYTES_PER_LAYER = (284230656 + 3859712) / 25


def run_e2e(args, lib, dev, rings, cases, step_bytes):
    """The sweep through the host-buffer C-ABI call: per case one tb200_gemv_stream_host (H2D of all activations of
    the ring, quantiser + persistent GEMV, D2H of all outputs)."""
    import torch
    from ternary_vllm_b200 import _lib as L
    stream = torch.cuda.current_stream(dev)
    R = {c: rings[c[:2]].R for c in cases}
    hx = {c: torch.randn(R[c], c[2], c[0], dtype=torch.bfloat16).pin_memory() for c in cases}
    hy = {c: torch.empty(R[c], c[2], c[1], dtype=torch.bfloat16).pin_memory() for c in cases}
    dx = {c: torch.empty(R[c], c[2], c[0], dtype=torch.bfloat16, device=dev) for c in cases}
    dy = {c: torch.empty(R[c], c[2], c[1], dtype=torch.bfloat16, device=dev) for c in cases}
    h2d = sum(hx[c].numel() * 2 for c in cases)
    d2h = sum(hy[c].numel() * 2 for c in cases)

    # Scheduling of the copies (PCIe is full duplex; the copies are 124 MB per step against 0.8 ms of kernels, so the step is
    # bound by the 80 MB of device -> host traffic): a case whose output exceeds E2E_PIECE_BYTES goes through the call in
    # several pieces of consecutive layers (the API takes any `count`), so that one piece's D2H copy overlaps the next
    # piece's kernels, and the jobs are issued in flow-shop (Johnson) order -- short uploads first, short downloads last --
    # round-robin over four streams.
    piece = int(os.environ.get("TB200_E2E_PIECE_BYTES", 12 << 20))
    jobs = []
    for c in cases:
        K, N, M = c
        out_bytes = R[c] * M * N * 2
        npieces = max(1, min(R[c], -(-out_bytes // piece))) if piece > 0 else 1
        per = -(-R[c] // npieces)
        for l0 in range(0, R[c], per):
            cnt = min(per, R[c] - l0)
            jobs.append({"c": c, "l0": l0, "cnt": cnt, "a": cnt * M * K * 2, "b": cnt * M * N * 2,
                         "ws": torch.empty(lib.tb200_gemv_stream_workspace(M, K, cnt), dtype=torch.uint8, device=dev)})
    first = sorted((j for j in jobs if j["a"] < j["b"]), key=lambda j: j["a"])
    last = sorted((j for j in jobs if j["a"] >= j["b"]), key=lambda j: -j["b"])
    jobs = first + last
    side = [torch.cuda.Stream(dev) for _ in range(4)]
    for i, j in enumerate(jobs):
        K, N, M = j["c"]
        ring, l0 = rings[(K, N)], j["l0"]
        j["args"] = (hx[j["c"]][l0].data_ptr(), L.TB200_BF16, M, K, ring.packed[l0].data_ptr(), ring.packed.stride(1),
                     ring.packed.stride(0), N, ring.scale[l0].data_ptr(), ring.bias[l0].data_ptr(), N, 100.0,
                     hy[j["c"]][l0].data_ptr(), L.TB200_BF16, j["cnt"], dx[j["c"]][l0].data_ptr(), dy[j["c"]][l0].data_ptr(),
                     j["ws"].data_ptr(), side[i % len(side)].cuda_stream)

    def step():
        for j in jobs:
            rc = lib.tb200_gemv_stream_host(*j["args"])
            if rc:
                L.check(rc, "gemv_stream_host")
        for st in side:
            st.synchronize()          # the caller reads hy after every step

    for _ in range(3):
        step()
    # the outputs this timing is about: the host-buffer call must reproduce the device-resident call bit for bit
    from ternary_vllm_b200 import ops
    for c in (cases[0], cases[-1]):
        ring = rings[c[:2]]
        want = ops.gemv_stream(hx[c].to(dev), ring.packed, ring.scale, ring.bias, 100.0)
        torch.cuda.synchronize()
        if not torch.equal(want.cpu(), hy[c]):
            raise RuntimeError(f"e2e: tb200_gemv_stream_host output differs from the device-resident call for case {c}")
    steps = max(3, min(args.steps, 20))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3          # host clock around synchronous steps: every step ends with the result on the host
    # what the PCIe link alone takes for the step's copies (same buffers, no kernels; uploads and downloads on two streams)
    def copies():
        with torch.cuda.stream(side[0]):
            for c in cases:
                dx[c].copy_(hx[c], non_blocking=True)
        with torch.cuda.stream(side[1]):
            for c in cases:
                hy[c].copy_(dy[c], non_blocking=True)
        side[0].synchronize(); side[1].synchronize()
    copies()
    t0 = time.perf_counter()
    for _ in range(5):
        copies()
    copies_ms = (time.perf_counter() - t0) * 1e3 / 5
    return {"value": step_bytes * steps / (ms * 1e-3) / 1e9, "unit": "GB/s", "h2d_bytes_per_step": h2d, "copies_only_ms_per_step": copies_ms,
            "d2h_bytes_per_step": d2h, "steps": steps, "ms_per_step": ms / steps,
            "calls_per_step": len(jobs),
            "api": "tb200_gemv_stream_host per sweep case (pinned host x -> device, quantiser + persistent GEMV over the "
                   "ring, device -> pinned host y); cases with more than 12 MB of output go in pieces of consecutive layers, "
                   "calls in flow-shop order round-robin on four streams, all synchronised at the end of every step"}


VIT_SHAPES = [(1024, 3072), (1024, 1024), (1024, 4096), (4096, 1024)]     # configs[3]: ViT-L/14 QKV, proj, fc1, fc2; M = 32 x 257


def run_vit(dev):
    """configs[3]: ViT-L/14 ternary linears, 32 images x 257 tokens.  One layer = tb200_linear_forward (activation
    quantiser + tcgen05 GEMM), replayed as a CUDA graph (the host time between the two eager launches is longer than the
    kernels), L2 flushed before every timed replay."""
    import torch
    from ternary_vllm_b200 import ops
    M = 32 * 257
    g = torch.Generator(device=dev).manual_seed(11)
    flush = torch.empty(int(256e6), dtype=torch.uint8, device=dev)
    st = torch.cuda.Stream(dev)
    rows = []
    for (K, N) in VIT_SHAPES:
        packed = ops.quantize_pack(torch.randn(N, K, generator=g, device=dev).clip_(-1, 1), "ternary")
        x = torch.randn(M, K, generator=g, device=dev, dtype=torch.bfloat16)
        scale = torch.full((N,), 1.0 / math.sqrt(K), device=dev)
        bias = torch.randn(N, generator=g, device=dev)
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        for _ in range(3):
            ops.ternary_matmul(x, packed, scale, bias, 100.0, out=out)
        torch.cuda.synchronize()
        with torch.cuda.stream(st):
            gl = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gl, stream=st):
                ops.ternary_matmul(x, packed, scale, bias, 100.0, out=out)
        torch.cuda.synchronize()
        ts = []
        for _ in range(10):
            flush.zero_()
            torch.cuda.synchronize()
            with torch.cuda.stream(st):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(st)
                gl.replay()
                e1.record(st)
                st.synchronize()
            ts.append(e0.elapsed_time(e1))
        t = statistics.median(ts) * 1e-3
        rows.append({"K": K, "N": N, "M": M, "us": t * 1e6, "tops_incl_act_quant": 2.0 * M * N * K / t / 1e12,
                     "tokens_per_s": M / t, "l2_policy": "L2 flushed (256 MB memset) before every timed replay"})
    return rows


def _timed_flush(fn, flush, reps=10):
    """Median device time of fn() with L2 flushed (256 MB memset) before every timed call."""
    import torch
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e-3)
    return statistics.median(ts)


def run_prep(dev, pk):
    """north_star subsystem (1) and (2) as standalone kernels: weight prep (quantize_pack, ternary_init, clip_latent) on a
    4096 x 11008 fp32 latent matrix and the per-token int8 activation quantiser, in GB/s of ALGORITHMIC bytes against the
    measured HBM rate.  ternary_init streams the matrix 12 times (mean, 10 thresholded means, rescale) and writes it once;
    the matrix (180 MB) exceeds L2, so every pass is an HBM pass."""
    import torch
    from ternary_vllm_b200 import ops
    g = torch.Generator(device=dev).manual_seed(21)
    N, K = 11008, 4096
    W = torch.randn(N, K, generator=g, device=dev) * 0.02
    flush = torch.empty(int(256e6), dtype=torch.uint8, device=dev)
    rows = []

    def add(name, fn, nbytes, note=None):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        t = _timed_flush(fn, flush)
        r = {"kernel": name, "us": t * 1e6, "algorithmic_bytes": nbytes, "gbs": nbytes / t / 1e9,
             "frac_hbm_measured": nbytes / t / 1e9 / pk["hbm_gbs"], "frac_hbm_spec_8TBs": nbytes / t / 1e9 / 8000.0}
        if note:
            r["note"] = note
        rows.append(r)
    Wc = W.clip(-1, 1)
    add("quantize_pack fp32 4096x11008 (tb200_quantize_pack)", lambda: ops.quantize_pack(Wc, "ternary"), N * K * 4 + N * K // 4)
    Wb = Wc.bfloat16()
    add("quantize_pack bf16 4096x11008", lambda: ops.quantize_pack(Wb, "ternary"), N * K * 2 + N * K // 4)
    Wl = (W * 60).contiguous()
    add("clip_latent fp32 4096x11008 (tb200_clip_latent)", lambda: ops.clip_latent_(Wl), 2 * N * K * 4)
    add("ternary_init uniform fp32 4096x11008 (tb200_ternary_init)", lambda: ops.ternary_init(W, per_row=False), 13 * N * K * 4,
        "12 reads + 1 write of the matrix (11 reductions with a grid-wide dependency between them)")
    add("ternary_init per-row fp32 4096x11008", lambda: ops.ternary_init(W, per_row=True), 2 * N * K * 4,
        "one CTA keeps a row on chip for all 11 passes: 1 read + 1 write")
    for (M, Kx) in ((4096, 4096), (8224, 1024), (8224, 4096)):
        x = torch.randn(M, Kx, generator=g, device=dev, dtype=torch.bfloat16)
        add(f"quant_act_i8 bf16 {M}x{Kx} (tb200_quant_act_i8, kperm)", lambda: ops.quant_act_i8(x, kperm=True), M * Kx * 3 + M * 8)
    return rows


def run_mlp(dev, pk):
    """DEPENDENT layers through the module API (TernaryMLP / ChainedLinears -> tb200_chain_forward): a
    stack of MLP blocks d -> f -> gelu_new -> d whose packed weights exceed L2, (a) as ONE chain call over the whole
    stack, (b) as one call per MLP block (what a decoder does: attention / residual / norm sit between the blocks), calls
    joined by programmatic dependent launch, (c) as one launch per layer (the round-1 `chain`) -- all as CUDA graphs."""
    import torch
    import ternary_vllm_b200 as tb
    from ternary_vllm_b200 import ops
    st = torch.cuda.Stream(dev)
    rows = []

    def time_graph(fn, reps):
        fn()
        st.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=st):
            fn()
        for _ in range(3):
            gr.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        st.synchronize()
        e0.record(st)
        for _ in range(reps):
            gr.replay()
        e1.record(st)
        st.synchronize()
        return e0.elapsed_time(e1) * 1e-3 / reps
    with torch.cuda.stream(st), torch.no_grad():
        for (d, f, n) in ((4096, 11008, 16), (2048, 8192, 40)):
            torch.manual_seed(d + f)
            blocks = []
            for _ in range(n):
                with torch.device(dev):                             # parameters are born on the GPU
                    fc1, fc2 = tb.ScaledTernaryLinear(d, f), tb.ScaledTernaryLinear(f, d)
                fc1.scale.data.fill_(1.0 / math.sqrt(d))
                fc2.scale.data.fill_(1.0 / math.sqrt(f))
                for l in (fc1, fc2):
                    l.freeze(release_latent=True)                   # deployment form: packed weights only
                blocks.append(tb.TernaryMLP(fc1, fc2, "gelu_new").eval())
            whole = tb.ChainedLinears([l for b_ in blocks for l in b_.layers], [a for b_ in blocks for a in b_.activations]).eval()
            for m_ in blocks + [whole]:
                m_.chain_max_tokens = 8                              # measure the chain kernel at every M (the default routes M > 2 layer by layer)
            for M in (1, 8):
                x = torch.randn(M, d, device=dev, dtype=torch.bfloat16)
                nbytes = n * (algorithmic_bytes(d, f, M) + algorithmic_bytes(f, d, M))

                def per_block():
                    y = x
                    for b_ in blocks:
                        y = b_(y, pdl=True)
                    return y

                def per_layer():
                    y = x
                    for b_ in blocks:
                        y = torch.nn.functional.gelu(b_.fc1(y), approximate="tanh")
                        y = b_.fc2(y)
                    return y
                t1 = time_graph(lambda: whole(x), 20)
                t2 = time_graph(per_block, 10)
                t3 = time_graph(per_layer, 5)
                mk = lambda t: {"us_per_layer": t / (2 * n) * 1e6, "gbs": nbytes / t / 1e9, "frac_hbm_measured": nbytes / t / 1e9 / pk["hbm_gbs"],
                                "tokens_per_s_per_block": M / (t / n)}
                rows.append({"d": d, "f": f, "blocks": n, "layers": 2 * n, "M": M, "packed_MB": n * 2 * d * f / 4 / 1e6,
                             "one_chain_call": mk(t1), "call_per_mlp_block_pdl": mk(t2), "launch_per_layer": mk(t3),
                             "api": "ternary_vllm_b200.TernaryMLP / ChainedLinears (tb200_chain_forward)"})
            del blocks, whole
            torch.cuda.empty_cache()
    return rows


def run_binary_1bit(dev, pk):
    """a ring of binary (+-1) 4096 x 11008 layers (packed bytes > 2 x L2 in both formats) as a chain of
    per-layer calls, the 2-bit format C (tb200_gemv, PDL) next to the 1-bit format B1 (tb200_gemv_b1_fused, PDL):
    the same integer sums from half the weight bytes."""
    import torch
    from ternary_vllm_b200 import ops, _lib as L
    K, N, R = 4096, 11008, 96
    g = torch.Generator(device=dev).manual_seed(5)
    p2, p1 = [], []
    for _ in range(R):
        W = torch.randn(N, K, generator=g, device=dev)
        p2.append(ops.quantize_pack(W, "binary"))
        p1.append(ops.quantize_pack_b1(W))
    del W
    scale = torch.full((N,), 1.0 / math.sqrt(K), device=dev)
    bias = torch.randn(N, generator=g, device=dev)
    st = torch.cuda.Stream(dev)
    rows = []
    with torch.cuda.stream(st), torch.no_grad():
        for M in (1, 2):
            x = torch.randn(M, K, generator=g, device=dev, dtype=torch.bfloat16)
            out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)

            def two_bit():
                for r in range(R):
                    ops.gemv(x, p2[r], scale, bias, 100.0, pdl=True, out=out)

            def one_bit():
                for r in range(R):
                    L.check(L.lib().tb200_gemv_b1_fused(x.data_ptr(), L.TB200_BF16, M, K, K, p1[r].data_ptr(), p1[r].stride(0), N,
                                                        scale.data_ptr(), bias.data_ptr(), 100.0, 0, out.data_ptr(), L.TB200_BF16, N,
                                                        L.FLAG_PDL, st.cuda_stream), "gemv_b1_fused")
            res = {}
            for name, fn, wbytes in (("two_bit_format_c", two_bit, N * K / 4), ("one_bit_format_b1", one_bit, N * K / 8)):
                fn()
                st.synchronize()
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr, stream=st):
                    fn()
                for _ in range(3):
                    gr.replay()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                st.synchronize()
                e0.record(st)
                for _ in range(10):
                    gr.replay()
                e1.record(st)
                st.synchronize()
                t = e0.elapsed_time(e1) * 1e-3 / 10 / R
                b = wbytes + M * K * 2 + M * N * 2 + N * 8
                res[name] = {"us_per_layer": t * 1e6, "tokens_per_s": M / t, "algorithmic_bytes_per_layer": b, "gbs": b / t / 1e9,
                             "frac_hbm_measured": b / t / 1e9 / pk["hbm_gbs"]}
            res.update({"K": K, "N": N, "M": M, "layers": R, "api": "chain of per-layer calls under one CUDA graph, PDL between launches",
                        "speedup_one_bit": res["two_bit_format_c"]["us_per_layer"] / res["one_bit_format_b1"]["us_per_layer"]})
            rows.append(res)
    return rows


def int8_peak(seconds=2.0):
    """Measured tcgen05 kind::i8 issue-rate ceiling of this GPU (tools/microbench/umma_i8_rate.cu, built by build())."""
    import subprocess
    exe = os.path.join(ROOT, "tools", "microbench", "umma_i8_rate.bin")
    if not os.path.exists(exe):
        return None
    try:
        out = subprocess.run([exe, "--seconds", str(seconds)], capture_output=True, text=True, timeout=120)
        return json.loads(out.stdout.strip().splitlines()[-1])
    except Exception as e:          # the roofline then falls back to the spec figure, and says so
        return {"error": repr(e)}


def run_gemm(dev, pk, M=4096, K=4096, N=11008):
    import torch
    from ternary_vllm_b200 import ops
    g = torch.Generator(device=dev).manual_seed(7)
    W = torch.randn(N, K, generator=g, device=dev).clip_(-1, 1)
    packed = ops.quantize_pack(W, "ternary")
    del W
    x = torch.randn(M, K, generator=g, device=dev, dtype=torch.bfloat16)
    scale = torch.full((N,), 1.0 / math.sqrt(K), device=dev)
    bias = torch.randn(N, generator=g, device=dev)
    flush = torch.empty(int(256e6), dtype=torch.uint8, device=dev)
    q, s, sq = ops.quant_act_i8(x, kperm=True)
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    for _ in range(3):
        ops.gemm_i8(q, s, sq, K, packed, scale, bias, 100.0)
        ops.ternary_matmul(x, packed, scale, bias, 100.0, out=out)
    torch.cuda.synchronize()
    # the whole layer (quantise + GEMM = tb200_linear_forward) as a CUDA graph: the Python call overhead
    # between the two launches is longer than the quantisation kernel
    st = torch.cuda.Stream(dev)
    with torch.cuda.stream(st):
        gl = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gl, stream=st):
            ops.ternary_matmul(x, packed, scale, bias, 100.0, out=out)
    torch.cuda.synchronize()
    t_mm, t_all = [], []
    for _ in range(10):
        flush.zero_()                                   # L2 flush between timed iterations
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.gemm_i8(q, s, sq, K, packed, scale, bias, 100.0)
        e1.record()
        torch.cuda.synchronize()
        t_mm.append(e0.elapsed_time(e1))
        flush.zero_()
        torch.cuda.synchronize()
        with torch.cuda.stream(st):
            e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e2.record(st)
            gl.replay()
            e3.record(st)
            st.synchronize()
        t_all.append(e2.elapsed_time(e3))
    ops_ = 2.0 * M * N * K
    mm, al = statistics.median(t_mm) * 1e-3, statistics.median(t_all) * 1e-3
    tops = ops_ / mm / 1e12
    # sustained: >= 2 s of back-to-back launches with the clock sampler on (the burst figure above is ten isolated launches)
    sampler = ClockSampler(dev.index or 0)
    n_sus = max(50, int(2.2 / mm))
    torch.cuda.synchronize()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n_sus):
        ops.gemm_i8(q, s, sq, K, packed, scale, bias, 100.0, out=out)
    e1.record()
    torch.cuda.synchronize()
    sus_clocks = sampler.stop()
    t_sus = e0.elapsed_time(e1) * 1e-3
    tops_sus = ops_ * n_sus / t_sus / 1e12
    pkq = int8_peak()
    peak_burst = peak_sus = None
    if pkq and "ts_n192" in pkq:
        peak_burst = max(pkq[k]["tops_burst"] for k in ("ss_n256", "ss_n192", "ts_n192"))
        peak_sus = max(pkq[k]["tops_sustained"] for k in ("ss_n256", "ts_n192"))
    roof = {"bound": "tensor", "achieved": tops, "unit": "TOP/s",
            "peak": peak_burst or 4500.0, "frac": tops / (peak_burst or 4500.0),
            "peak_source": ("measured: tools/microbench/umma_i8_rate.cu, best burst rate of tcgen05.mma kind::i8 issued back to back on all SMs"
                            if peak_burst else "spec: dense int8 4.5 POPS (microbenchmark binary missing)"),
            "frac_of_spec_4500": tops / 4500.0, "frac_of_2x_measured_bf16": tops / (2 * pk["bf16_tflops"]),
            "traffic": GEMM_TRAFFIC_BYTES}
    return {"workload": f"configs[2] prefill GEMM M={M} K={K} N={N}, int8 activations, tcgen05 kind::i8 (weights through TMEM)",
            "ms_gemm": mm * 1e3, "ms_quant_plus_gemm": al * 1e3, "tops": tops, "tops_incl_act_quant": ops_ / al / 1e12,
            "tokens_per_s": M / al,
            "sustained": {"tops": tops_sus, "seconds": t_sus, "launches": n_sus, "clocks": sus_clocks,
                          "frac_of_measured_sustained_int8": (tops_sus / peak_sus) if peak_sus else None,
                          "frac_of_spec_4500": tops_sus / 4500.0},
            "int8_peak": pkq, "roofline": roof,
            "l2_policy": "L2 flushed (256 MB memset) between timed iterations (burst); none for the sustained run (operands 62 MB + 90 MB output)"}


# dram__bytes_read.sum + dram__bytes_write.sum of umma_ts_kernel<bf16,192,1> on M=4096 K=4096 N=11008 (profiles/r2_umma_gemm.summary.txt,
# 130.4 us under ncu, tensor pipe 77.8 % of active cycles): 28.2 MB read (int8 activations 16.8 MB + packed weights 11.3 MB) and
# 40.7 MB written of the 90.2 MB bf16 output -- the rest of the output was still in the 126 MB L2 when the kernel ended
GEMM_TRAFFIC_BYTES = 28248320 + 40744448


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-gemm", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the mlp_chain and prep blocks")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
