#!/usr/bin/env python
"""Skinny decode timing (SURVEY.md §8 rows a9-a12 at n <= 8): one weight matrix x T = 1..8 tokens through
tb200_mul_mat_acts (activations already quantized), Llama-3-8B shapes. Each shape is timed as a CUDA graph that walks
enough distinct copies of the matrix to exceed L2 twice, so every launch streams from HBM.
    python tools/bench_skinny.py [--types q4_0] [--tokens 1,2,3,4,8] [--iters 20]
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tools"))

from bench_prefill import TYPES, synth   # noqa: E402

SHAPES = [(4096, 4096), (14336, 4096), (4096, 14336), (128256, 4096)]
BPE = {"q4_0": 18 / 32, "q8_0": 34 / 32, "q4_K": 144 / 256, "q6_K": 210 / 256, "f16": 2.0}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--types", default="q4_0")
    ap.add_argument("--tokens", default="1,2,3,4,6,8")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--shape", default="", help="N,K: only this shape")
    a = ap.parse_args()
    import torch
    from llamafile_b200.build import build
    build()
    import llamafile_b200 as tb
    assert tb.link(), "no sm_100 device"
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(device=dev)
    tb.set_stream(stream.cuda_stream)
    for name in a.types.split(","):
        wtype = TYPES[name][0]
        for (N, K) in ([tuple(int(v) for v in a.shape.split(","))] if a.shape else SHAPES):
            wbytes = N * K * BPE[name]
            copies = max(2, int(2 * 126e6 * 2 // wbytes) + 1) if wbytes < 256e6 else 2
            copies = min(copies, 64)
            ws = [tb.Weights(synth(torch, dev, name, N, K, seed=i), wtype, N, K) for i in range(copies)]
            for T in [int(t) for t in a.tokens.split(",")]:
                x = torch.randn(T, K, device=dev)
                acts = tb.Acts(wtype, K, T)
                out = torch.empty(T, N, device=dev)
                with torch.cuda.stream(stream):
                    acts.quantize(x, T)
                    for w in ws[:2]:
                        tb.mul_mat_acts(w, acts, out)
                    stream.synchronize()
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g, stream=stream):
                        for w in ws:
                            tb.mul_mat_acts(w, acts, out)
                    for _ in range(3):
                        g.replay()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(stream)
                    for _ in range(a.iters):
                        g.replay()
                    e1.record(stream)
                    stream.synchronize()
                us = e0.elapsed_time(e1) * 1e3 / a.iters / copies
                byts = wbytes + T * K * 4 + T * N * 4
                print(json.dumps({"op": "mul_mat", "type": name, "N": N, "K": K, "tokens": T, "copies": copies, "us_per_launch": us,
                                  "GBps": byts / (us * 1e-6) / 1e9, "TFLOPS": 2.0 * T * N * K / (us * 1e-6) / 1e12}), flush=True)
                del g
                acts.free() if hasattr(acts, "free") else None
            for w in ws:
                w.free()


if __name__ == "__main__":
    main()
