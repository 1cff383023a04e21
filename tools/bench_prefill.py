#!/usr/bin/env python
"""Prefill GEMM timing (BASELINE configs[2]): Q8_0 / Q4_0 (and Q4_K, F16) weights x T tokens on tcgen05.
TFLOPS = 2*T*N*K / CUDA-event time of tb200_mul_mat (activation quantize + prep + GEMM) and of the GEMM launch alone.
    python tools/bench_prefill.py [--types q8_0,q4_0] [--tokens 512,4096] [--iters 20]
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

TYPES = {"q4_0": (2, 18, 32), "q8_0": (8, 34, 32), "q4_K": (12, 144, 256), "f16": (1, 2, 1)}
SHAPES = [(4096, 4096), (14336, 4096), (4096, 14336)]


def synth(torch, dev, name, N, K, seed=0):
    t, ts, blck = TYPES[name]
    g = torch.Generator(device=dev); g.manual_seed(seed)
    if name == "f16":
        return (torch.randn(N, K, device=dev, generator=g) * 0.02).to(torch.float16).view(torch.uint8).reshape(-1)
    nb = N * K // blck
    raw = torch.randint(0, 256, (nb, ts), dtype=torch.uint8, device=dev, generator=g)
    d = (torch.rand(nb, device=dev, generator=g) + 0.5) * 2e-3
    raw[:, 0:2] = d.to(torch.float16).view(torch.uint8).reshape(nb, 2)
    if name == "q4_K":
        raw[:, 2:4] = (d * 2).to(torch.float16).view(torch.uint8).reshape(nb, 2)
    return raw.reshape(-1)


def run(types, tokens, shapes=SHAPES, iters=20, warmup=3):
    import torch
    from llamafile_b200.build import build
    build()
    import llamafile_b200 as tb
    assert tb.link(), "no sm_100 device"
    dev = torch.device("cuda", torch.cuda.current_device())
    stream = torch.cuda.Stream(device=dev)
    tb.set_stream(stream.cuda_stream)
    res = []
    for name in types:
        for (N, K) in shapes:
            w = tb.Weights(synth(torch, dev, name, N, K), TYPES[name][0], N, K)
            for T in tokens:
                x = torch.randn(T, K, device=dev)
                out = torch.empty(T, N, device=dev)
                flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > L2: weights come from HBM every iteration
                with torch.cuda.stream(stream):
                    for _ in range(warmup):
                        tb.mul_mat(w, x, T, out)
                    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
                    for a, b in ev:
                        flush.zero_()
                        a.record(stream)
                        tb.mul_mat(w, x, T, out)
                        b.record(stream)
                    stream.synchronize()
                ms = sorted(a.elapsed_time(b) for a, b in ev)
                med = ms[len(ms) // 2]
                fl = 2.0 * T * N * K
                res.append({"type": name, "N": N, "K": K, "T": T, "ms_median": med, "ms_min": ms[0], "TFLOPS": fl / (med * 1e-3) / 1e12,
                            "TFLOPS_best": fl / (ms[0] * 1e-3) / 1e12, "includes": "activation quantize + fp16 prep + tcgen05 GEMM",
                            "l2": "256 MiB flush between iterations"})
            w.free()
    tb.set_stream(0)
    return res


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--types", default="q8_0,q4_0")
    ap.add_argument("--tokens", default="512,4096")
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    for r in run(a.types.split(","), [int(t) for t in a.tokens.split(",")], iters=a.iters):
        print(json.dumps(r), flush=True)
