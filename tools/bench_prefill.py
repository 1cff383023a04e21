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

TYPES = {"q4_0": (2, 18, 32), "q8_0": (8, 34, 32), "q4_K": (12, 144, 256), "q6_K": (14, 210, 256), "f16": (1, 2, 1)}
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
    if name == "q6_K":   # d sits in the last two bytes of block_q6_K
        raw[:, 0:2] = torch.randint(0, 256, (nb, 2), dtype=torch.uint8, device=dev, generator=g)
        raw[:, 208:210] = (d * 1e-2).to(torch.float16).view(torch.uint8).reshape(nb, 2)
    return raw.reshape(-1)


def cpu_reference(name, N, K, T, budget_s=4.0):
    """The reference's own CPU tinyBLAS/iqk path (oracle/_ref) on a bounded slice of the same problem (all host threads)."""
    import time
    import numpy as np
    from oracle import binding as OB
    import os
    ref = OB.reference()
    if ref is None:
        raise RuntimeError("oracle/_ref not built")
    nth = os.cpu_count() or 1
    t = TYPES[name][0]
    Tn, Nn = min(T, 256), min(N, 2048)     # bounded sample: Tn tokens x Nn weight rows, full K
    rng = np.random.default_rng(0)
    if name == "f16":
        W = (rng.standard_normal((Nn, K)) * 0.02).astype(np.float16).view(np.uint8).reshape(-1)
    else:
        ts, blck = TYPES[name][1], TYPES[name][2]
        W = rng.integers(0, 256, Nn * K // blck * ts, dtype=np.uint8)
        Wb = W.reshape(-1, ts)
        Wb[:, 0:2] = np.full(Wb.shape[0], 1e-3, np.float16).view(np.uint8).reshape(-1, 2)
        if name == "q4_K":
            Wb[:, 2:4] = np.full(Wb.shape[0], 2e-3, np.float16).view(np.uint8).reshape(-1, 2)
    X = rng.standard_normal((Tn, K)).astype(np.float32)
    ref.mul_mat(W, t, X, nth=nth)
    n, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < budget_s or n < 2:
        ref.mul_mat(W, t, X, nth=nth)
        n += 1
    dt = (time.perf_counter() - t0) / n
    return {"value": 2.0 * Tn * Nn * K / dt / 1e12, "unit": "TFLOPS", "cores": nth, "kind": "reference",
            "sample": f"{Tn} tokens x {Nn} rows x K={K}, {n} reps, activation quantization included, ISA variant {ref.variant}"}


def run(types, tokens, shapes=SHAPES, iters=20, warmup=3, cpu=False):
    import os
    import torch
    import torch.distributed as dist
    from llamafile_b200.build import build
    from llamafile_b200.shard import shard_rows
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG", "WARN")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    build()
    import llamafile_b200 as tb
    assert tb.link(), "no sm_100 device"
    dev = torch.device("cuda", torch.cuda.current_device())
    stream = torch.cuda.Stream(device=dev)
    tb.set_stream(stream.cuda_stream)
    res = []
    if world > 1:
        return run_sharded(torch, dist, tb, dev, stream, shard_rows, rank, world, types, tokens, shapes, iters, warmup)
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
                if cpu:
                    try:
                        res[-1]["cpu_baseline"] = cpu_reference(name, N, K, T)
                    except Exception as e:   # noqa: BLE001
                        res[-1]["cpu_baseline"] = {"value": None, "sample": f"failed: {e}"}
                print("#", name, N, K, T, round(med * 1000, 1), "us", file=sys.stderr, flush=True)
            w.free()
    tb.set_stream(0)
    return res


def run_sharded(torch, dist, tb, dev, stream, shard_rows, rank, world, types, tokens, shapes, iters, warmup):
    """Row-sharded prefill: every rank multiplies its N/g weight rows by the replicated activations. Exchange:
    (a) fused — the GEMM epilogue stores each output segment into every rank's symmetric [T, N] buffer over NVLink, then one
    flag barrier; (b) ncclAllGather of the [T, N/g] slices on the same stream. Device time, max over ranks."""
    from llamafile_b200.peer import PeerGroup
    res = []
    for name in types:
        for (N, K) in shapes:
            lo, hi = shard_rows(N, rank, world)
            n_loc = hi - lo
            n_max = max(b - a for a, b in (shard_rows(N, r, world) for r in range(world)))
            w = tb.Weights(synth(torch, dev, name, n_loc, K, seed=rank), TYPES[name][0], n_loc, K)
            for T in tokens:
                x = torch.randn(T, K, device=dev)
                out = torch.zeros(T, n_max, device=dev)          # padded to the widest shard so the gather is regular
                gathered = torch.empty(world, T, n_max, device=dev)
                flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

                def timed(with_gather):
                    ms = []
                    with torch.cuda.stream(stream):
                        for _ in range(warmup):
                            tb.mul_mat(w, x, T, out, ldo=n_max)
                            if with_gather:
                                dist.all_gather_into_tensor(gathered, out)
                        for _ in range(iters):
                            flush.zero_()
                            dist.barrier()
                            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                            e0.record(stream)
                            tb.mul_mat(w, x, T, out, ldo=n_max)
                            if with_gather:
                                dist.all_gather_into_tensor(gathered, out)
                            e1.record(stream)
                            stream.synchronize()
                            t = torch.tensor([e0.elapsed_time(e1)], device=dev)
                            dist.all_reduce(t, op=dist.ReduceOp.MAX)
                            ms.append(float(t.item()))
                    ms.sort()
                    return ms[len(ms) // 2]
                m_gemm, m_all = timed(False), timed(True)
                # fused exchange
                pg = PeerGroup(world, rank, T * N * 4)
                acts = tb.Acts(TYPES[name][0], K, T)
                ms = []
                with torch.cuda.stream(stream):
                    for it in range(warmup + iters):
                        flush.zero_()
                        dist.barrier()
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record(stream)
                        acts.quantize(x, T)
                        pg.mul_mat_gather([w], acts, [0], [N], [lo])
                        pg.barrier()
                        e1.record(stream)
                        stream.synchronize()
                        tt = torch.tensor([e0.elapsed_time(e1)], device=dev)
                        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                        if it >= warmup:
                            ms.append(float(tt.item()))
                ms.sort()
                m_fused = ms[len(ms) // 2]
                assert not pg.timed_out(), "peer barrier timed out"
                acts.free()
                fl = 2.0 * T * N * K
                if rank == 0:
                    res.append({"type": name, "N": N, "K": K, "T": T, "n_gpus": world, "ms_gemm": m_gemm, "ms_gemm_plus_allgather": m_all,
                                "ms_fused_gather": m_fused, "TFLOPS_fused_gather": fl / (m_fused * 1e-3) / 1e12,
                                "TFLOPS": fl / (m_all * 1e-3) / 1e12, "TFLOPS_gemm_only": fl / (m_gemm * 1e-3) / 1e12,
                                "allgather_bytes_per_rank": T * n_max * 4, "parallelism": f"row-shard x{world} + ncclAllGather",
                                "includes": "activation quantize + fp16 prep + tcgen05 GEMM (+ all-gather)", "l2": "256 MiB flush between iterations"})
            w.free()
    dist.barrier()
    return res


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--types", default="q8_0,q4_0")
    ap.add_argument("--tokens", default="512,4096")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--shapes", default="", help="N:K,N:K,... (default: Llama-3-8B layer shapes); 70b = Llama-3-70B shapes")
    ap.add_argument("--cpu", action="store_true", help="time the reference's CPU path on a bounded slice next to each line")
    a = ap.parse_args()
    shapes = SHAPES
    if a.shapes == "70b":
        shapes = [(8192, 8192), (28672, 8192), (8192, 28672)]
    elif a.shapes:
        shapes = [tuple(int(v) for v in sh.split(":")) for sh in a.shapes.split(",")]
    for r in run(a.types.split(","), [int(t) for t in a.tokens.split(",")], shapes=shapes, iters=a.iters, cpu=a.cpu):
        print(json.dumps(r), flush=True)
    import os as _os
    import sys as _sys
    _sys.stdout.flush()
    _os._exit(0)   # process-group teardown can stall; everything is printed
