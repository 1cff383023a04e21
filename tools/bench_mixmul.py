#!/usr/bin/env python
"""mul_mat_id timing (BASELINE configs[3]): Mixtral-8x7B expert matrices (8 experts, top-2 routing) through
tb200_mixmul_dev — device-side expert grouping + grouped decode GEMV (small batches) or grouped tcgen05 GEMM.
Reports per (matrix, batch): time, algorithmic GB/s (touched experts' weights + activations + outputs) and TFLOPS.
    python tools/bench_mixmul.py [--type q4_0] [--batches 1,8,64,512] [--iters 20]
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tools"))

from bench_prefill import TYPES, synth   # noqa: E402

EXPERTS, TOP = 8, 2
# Mixtral-8x7B: hidden 4096, expert FFN 14336: w1/w3 (gate/up) [14336 x 4096], w2 (down) [4096 x 14336]
SHAPES = [("ffn_gate_exps", 14336, 4096), ("ffn_down_exps", 4096, 14336)]
BPE = {"q4_0": 18 / 32, "q8_0": 34 / 32, "f16": 2.0}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--type", default="q4_0")
    ap.add_argument("--batches", default="1,8,64,512")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--graph", type=int, default=4, help="also time a CUDA graph of this many calls over as many distinct expert stacks (0 = off)")
    a = ap.parse_args()
    import torch
    from llamafile_b200.build import build
    build()
    import llamafile_b200 as tb
    assert tb.link(), "no sm_100 device"
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(device=dev)
    tb.set_stream(stream.cuda_stream)
    g = torch.Generator(device=dev); g.manual_seed(1)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for name in a.type.split(","):
        for (mat, rows, cols) in SHAPES:
            w = tb.Weights(synth(torch, dev, name, rows * EXPERTS, cols), TYPES[name][0], rows * EXPERTS, cols)
            # graph mode: a decode step calls mul_mat_id once per layer, each on its own expert stack, from one captured graph; the
            # distinct stacks (a.graph x 2-8 touched experts x 33 MB) exceed L2, and launch overhead of the eager loop is gone
            more = [tb.Weights(synth(torch, dev, name, rows * EXPERTS, cols), TYPES[name][0], rows * EXPERTS, cols) for _ in range(max(0, a.graph - 1))]
            for tokens in [int(b) for b in a.batches.split(",")]:
                x = torch.randn(tokens, 1, cols, device=dev, generator=g)
                # top-2 routing: two distinct experts per token, uniformly random
                plan = torch.stack([torch.randperm(EXPERTS, device=dev, generator=g)[:TOP] for _ in range(tokens)]).to(torch.int32).contiguous()
                out = torch.empty(tokens, TOP, rows, device=dev)
                touched = int(torch.unique(plan).numel())
                with torch.cuda.stream(stream):
                    for _ in range(3):
                        assert tb.mixmul(w, EXPERTS, x, 1, plan, TOP, tokens, out)
                    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.iters)]
                    for e0, e1 in ev:
                        flush.zero_()
                        e0.record(stream)
                        tb.mixmul(w, EXPERTS, x, 1, plan, TOP, tokens, out)
                        e1.record(stream)
                    stream.synchronize()
                ms = sorted(e0.elapsed_time(e1) for e0, e1 in ev)
                med = ms[len(ms) // 2]
                graph_ms = None
                if a.graph:
                    stacks = [w] + more
                    plans = [torch.stack([torch.randperm(EXPERTS, device=dev, generator=g)[:TOP] for _ in range(tokens)]).to(torch.int32).contiguous()
                             for _ in stacks]
                    touched_g = sum(int(torch.unique(p).numel()) for p in plans)
                    with torch.cuda.stream(stream):
                        for ws, pl in zip(stacks, plans):
                            tb.mixmul(ws, EXPERTS, x, 1, pl, TOP, tokens, out)
                        stream.synchronize()
                        gr = torch.cuda.CUDAGraph()
                        with torch.cuda.graph(gr, stream=stream):
                            for ws, pl in zip(stacks, plans):
                                tb.mixmul(ws, EXPERTS, x, 1, pl, TOP, tokens, out)
                        for _ in range(3):
                            gr.replay()
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record(stream)
                        for _ in range(a.iters):
                            gr.replay()
                        e1.record(stream)
                        stream.synchronize()
                    graph_ms = e0.elapsed_time(e1) / a.iters / len(stacks)
                    gbytes = (touched_g * rows * cols * BPE[name]) / len(stacks) + tokens * cols * 4 + tokens * TOP * rows * 4
                    del gr
                byts = touched * rows * cols * BPE[name] + tokens * cols * 4 + tokens * TOP * rows * 4
                fl = 2.0 * tokens * TOP * rows * cols
                print(json.dumps({"op": "mul_mat_id", "type": name, "matrix": mat, "rows": rows, "cols": cols, "experts": EXPERTS, "top": TOP,
                                  "tokens": tokens, "experts_touched": touched, "ms_median": med, "ms_min": ms[0],
                                  "GBps": byts / (med * 1e-3) / 1e9, "TFLOPS": fl / (med * 1e-3) / 1e12,
                                  "graph_calls": a.graph, "graph_ms_per_call": graph_ms,
                                  "graph_GBps": gbytes / (graph_ms * 1e-3) / 1e9 if graph_ms else None,
                                  "graph_TFLOPS": fl / (graph_ms * 1e-3) / 1e12 if graph_ms else None,
                                  "includes": "activation quantize + device-side grouping + grouped kernel", "l2": "256 MiB flush between iterations"}), flush=True)
            w.free()
            for m in more:
                m.free()


if __name__ == "__main__":
    main()
