#!/usr/bin/env python
"""Where does a decode-plan group spend its time? Runs a plan of identical groups and prints per-phase SM-clock deltas
(tb200_plan_trace) averaged over groups, for a few CTAs, plus the CUDA-event time per group.
    python tools/plan_probe.py [--n 4096 --k 4096 --groups 32]
The stamps exist only in a probe build of the library:
    tools/build_variant.sh plan_dbg tb200_plan.cu -DTB200_PLAN_DEBUG=1 [-DTB200_PLAN_TRACE_ITEMS]
    TB200_LIB=tools/_variants/plan_dbg.so python tools/plan_probe.py ..."""
import argparse
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def lib_has_no_trace(plan):
    from llamafile_b200.lib import lib
    return lib().tb200_plan_trace(plan.handle, None, 0) == -2


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=4096)
    ap.add_argument("--k", type=int, default=4096)
    ap.add_argument("--groups", type=int, default=32)
    ap.add_argument("--mats", type=int, default=1)
    a = ap.parse_args()
    import torch
    from llamafile_b200.build import build
    build()
    import llamafile_b200 as tb
    assert tb.link()
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(device=dev)
    tb.set_stream(stream.cuda_stream)
    g = torch.Generator(device=dev); g.manual_seed(1)
    groups, keep = [], []
    for gi in range(a.groups):
        ws, outs = [], []
        for m in range(a.mats):
            nb = a.n * a.k // 32
            raw = torch.randint(0, 256, (nb, 18), dtype=torch.uint8, device=dev, generator=g)
            raw[:, 0:2] = torch.full((nb,), 2e-3, device=dev).to(torch.float16).view(torch.uint8).reshape(nb, 2)
            ws.append(tb.Weights(raw.reshape(-1), 2, a.n, a.k))
            outs.append(torch.zeros(a.n, device=dev))
        x = torch.randn(a.k, device=dev)
        groups.append({"w": ws, "out": outs, "x": x})
        keep.append((ws, outs, x))
    plan = tb.Plan(groups)
    with torch.cuda.stream(stream):
        for _ in range(3):
            plan.run()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(10):
            plan.run()
        e1.record(stream)
        stream.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 10 / a.groups
    byts = a.mats * a.n * a.k * 18 / 32
    print(f"N={a.n} K={a.k} mats={a.mats}: {us:.2f} us per group, {byts / us / 1e3:.0f} GB/s, status {plan.status()}")
    if plan.trace() is None and lib_has_no_trace(plan):
        print('  (no phase stamps in this build: see the docstring)')
        return
    with torch.cuda.stream(stream):
        plan.run()
        stream.synchronize()
    t = plan.trace().astype(np.float64)
    names = ["warp0 loop top -> CTA arrived", "fence+atomic", "poll", "release -> warp0", "stage acts", "compute (warp 0)"]
    d = np.diff(t[:, :, :7], axis=2)          # [groups, ctas, 6]
    d = d[2:]                                  # skip the first groups (cold)
    for i, nm in enumerate(names):
        col = d[:, :, i]
        print(f"  {nm:26s} mean {col.mean():8.0f}  median {np.median(col):8.0f}  p95 {np.percentile(col, 95):8.0f}  max {col.max():8.0f} cycles")
    t = t[2:]
    has = t[:, :, 8] > 0                      # CTAs whose warp 0 owned a pair in the group
    def rep(nm, a):
        a = a[has]
        print(f"  {nm:34s} mean {a.mean():8.0f}  median {np.median(a):8.0f}  p95 {np.percentile(a, 95):8.0f}")
    rep("sync warp: release fence", t[:, :, 13] - t[:, :, 1])
    rep("sync warp: atomic add round trip", t[:, :, 2] - t[:, :, 13])
    rep("release -> x loads landed", t[:, :, 12] - t[:, :, 4])
    rep("x landed -> quantized (warp 0)", t[:, :, 7] - t[:, :, 12])
    rep("quantized -> CTA barrier passed", t[:, :, 5] - t[:, :, 7])
    rep("staged -> first item landed", t[:, :, 8] - t[:, :, 5])
    rep("first item: block dots", t[:, :, 9] - t[:, :, 8])
    rep("first item: refill + advance", t[:, :, 10] - t[:, :, 9])
    rep("   expect_tx + syncwarp", t[:, :, 14] - t[:, :, 9])
    rep("   bulk copy issue", t[:, :, 15] - t[:, :, 14])
    rep("   advance + prepare", t[:, :, 10] - t[:, :, 15])
    rep("first pair: rest + reduce + store", t[:, :, 11] - t[:, :, 10])
    items = np.maximum(t[:, :, 20], 1)
    print(f"  warp 0, all its items of a group: {t[:, :, 20][has].mean():.2f} items; per item: wait {(t[:, :, 16] / items)[has].mean():.0f}  dots {(t[:, :, 17] / items)[has].mean():.0f}"
          f"  refill {(t[:, :, 18] / items)[has].mean():.0f}  reduce+store (per unit) {(t[:, :, 19] / items)[has].mean():.0f} cycles")
    per_group = t[1:, :, 0] - t[:-1, :, 0]
    print(f"  loop-top to loop-top        mean {per_group.mean():8.0f} cycles")
    # arrival skew: spread of the arrival stamps (tr[2]) across CTAs of one group is not comparable across SMs (per-SM clocks),
    # so report the spread of 'poll' instead: the CTA that arrives last polls least
    print(f"  poll min over CTAs (last arriver)  mean {d[:, :, 2].min(axis=1).mean():8.0f}")


if __name__ == "__main__":
    main()
