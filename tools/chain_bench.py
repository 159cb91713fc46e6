"""Timing of tb200_chain_forward on stacks of dependent MLP blocks (device time, CUDA-graph replay, weights > L2).

    python tools/chain_bench.py [--mlps 32]
"""
import argparse
import json
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from ternary_vllm_b200 import ops


def alg_bytes(K, N, M):
    return N * K / 4 + M * K * 2 + M * N * 2 + N * 8


def build_stack(d, f, n_mlps, dev, seed):
    g = torch.Generator(device=dev).manual_seed(seed)
    layers = []
    for _ in range(n_mlps):
        for (K, N, act) in ((d, f, "gelu_new"), (f, d, None)):
            W = torch.randn(N, K, generator=g, device=dev).clip_(-1, 1)
            layers.append((ops.quantize_pack(W, "ternary"), torch.full((N,), 1.0 / math.sqrt(K), device=dev),
                           torch.randn(N, generator=g, device=dev) * 0.1, 100.0, act))
    return layers


def time_graph(fn, stream, reps):
    fn()
    stream.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr, stream=stream):
        fn()
    for _ in range(3):
        gr.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stream.synchronize()
    e0.record(stream)
    for _ in range(reps):
        gr.replay()
    e1.record(stream)
    stream.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mlps", type=int, default=32)
    ap.add_argument("--fast", action="store_true", help="only the one-call form, 4096/11008 stack")
    ap.add_argument("--tokens", type=int, nargs="*", default=[1, 2, 4, 8])
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    st = torch.cuda.Stream(dev)
    rows = []
    with torch.cuda.stream(st):
        for (d, f, n) in ((4096, 11008, args.mlps), (2048, 8192, min(64, 2 * args.mlps)))[: 1 if args.fast else 2]:
            layers = build_stack(d, f, n, dev, d + f)
            for M in args.tokens:
                x = torch.randn(M, d, device=dev, dtype=torch.bfloat16)
                plan = ops.ChainPlan(layers, d, torch.bfloat16, max_tokens=M)
                out = torch.empty(M, d, device=dev, dtype=torch.bfloat16)
                t = time_graph(lambda: plan.forward(x, out=out), st, 20)
                b = n * (alg_bytes(d, f, M) + alg_bytes(f, d, M))
                if args.fast:
                    print(json.dumps({"d": d, "f": f, "M": M, "one_call_us_per_layer": t / (2 * n) * 1e6, "one_call_gbs": b / t / 1e9}), flush=True)
                    continue
                # the same stack as one chain call PER MLP block (2 layers), calls chained with PDL
                plans = [ops.ChainPlan(layers[2 * i: 2 * i + 2], d, torch.bfloat16, max_tokens=M) for i in range(n)]
                bufs = [torch.empty(M, d, device=dev, dtype=torch.bfloat16) for _ in range(2)]

                def per_block():
                    y = x
                    for i, pl in enumerate(plans):
                        y = pl.forward(y, out=bufs[i & 1], pdl=True)
                t2 = time_graph(per_block, st, 10)
                # and as a chain of per-layer launches (r1's `chain`): tb200_gemv + PDL, gelu as a torch op
                hb = [torch.empty(M, f, device=dev, dtype=torch.bfloat16), torch.empty(M, d, device=dev, dtype=torch.bfloat16)]

                def per_layer():
                    y = x
                    for i, (pw, sc, bi, clip, act) in enumerate(layers):
                        y = ops.gemv(y, pw, sc, bi, clip, pdl=True, out=hb[i & 1])
                        if act:
                            y = torch.nn.functional.gelu(y, approximate="tanh")
                t3 = time_graph(per_layer, st, 5)
                rows.append({"d": d, "f": f, "mlps": n, "M": M, "layers": 2 * n,
                             "one_call_us_per_layer": t / (2 * n) * 1e6, "one_call_gbs": b / t / 1e9,
                             "call_per_mlp_us_per_layer": t2 / (2 * n) * 1e6, "call_per_mlp_gbs": b / t2 / 1e9,
                             "launch_per_layer_us_per_layer": t3 / (2 * n) * 1e6, "launch_per_layer_gbs": b / t3 / 1e9})
                print(json.dumps(rows[-1]), flush=True)
                del plans, plan
            del layers
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
