#!/usr/bin/env python
"""bench.py — BASELINE.json's metric on B200: Q4_0 decode GEMV GB/s (+ Q8_0/Q4_0 prefill GEMM TFLOPS) at Llama-3-8B shapes.

A "step" is one decode token's worth of weight matmuls of Llama-3-8B (BASELINE configs[1]: M=1, K=4096,
N=4096/14336/128256 plus the model's other matrices): 32 layers x {wq, wk, wv, wo, gate, up, down} + the output
projection = 225 GEMVs in 129 launch groups (matrices sharing an input — wq/wk/wv, gate/up — form one group) over
4.22 GB of distinct Q4_0 weights (>> 126 MB L2: every step streams from HBM, no flush needed). Each group's fp32 input is
quantized to Q8_0 exactly as ggml does before llamafile_sgemm, inside the timed region.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload NAME] [--decode plan|oneshot]

value    : algorithmic GB/s of the whole step, inputs resident in HBM. Default (Q4_0, 1 GPU): the DECODE PLAN, the whole step
           as ONE persistent kernel (tb200_plan.cu: per-warp TMA rings, in-kernel quantization, group g waits for g-1).
           `--decode oneshot` (and the other weight types / N > 1): one quantizer + one GEMV launch per group from a CUDA graph.
roofline : the dominant kernel — the plan kernel (it IS the step), or the GEMV launches alone for oneshot — with CUDA events;
           plus roofline.prefill[] (tcgen05 GEMM, TFLOPS sustained for >= 1 s per point), roofline.mixmul[], roofline.quantizer.
e2e      : the same step through the reference-facing C entry point llamafile_sgemm() with HOST buffers.
cpu_baseline / --impl reference: the reference's own CPU tinyBLAS/iqk path (oracle/_ref) over the SAME full step on the host cores.
N > 1    : launch groups >= --split-min-mb row-sharded over the ranks, outputs exchanged by peer stores fused into the kernels.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

Q4_0, Q8_0, Q4_K, Q6_K, F16, F32 = 2, 8, 12, 14, 1, 0
BPE = {Q4_0: 18 / 32, Q8_0: 34 / 32, Q4_K: 144 / 256, Q6_K: 210 / 256, F16: 2.0}
TS_BLCK = {Q4_0: (18, 32), Q8_0: (34, 32), Q4_K: (144, 256), Q6_K: (210, 256)}
SPEC_HBM_GBS, SPEC_BF16_TFLOPS = 8000.0, 2250.0       # the figures north_star names

_LLAMA8B = [("wq", 4096, 4096, "attn"), ("wk", 1024, 4096, "attn"), ("wv", 1024, 4096, "attn"), ("wo", 4096, 4096, "o"),
            ("gate", 14336, 4096, "ffn"), ("up", 14336, 4096, "ffn"), ("down", 4096, 14336, "down")]
WORKLOADS = {
    # name: (weight type, layers, per-layer [(name, N, K, input-group)], [(name, N, K, group)] tail)
    "llama3-8b-q4_0-decode": (Q4_0, 32, _LLAMA8B, [("output", 128256, 4096, "out")]),
    "llama3-8b-q6_K-decode": (Q6_K, 32, _LLAMA8B, [("output", 128256, 4096, "out")]),
    "llama3-70b-q4_K-decode": (Q4_K, 80, [("wq", 8192, 8192, "attn"), ("wk", 1024, 8192, "attn"), ("wv", 1024, 8192, "attn"),
                                          ("wo", 8192, 8192, "o"), ("gate", 28672, 8192, "ffn"), ("up", 28672, 8192, "ffn"),
                                          ("down", 8192, 28672, "down")], [("output", 128256, 8192, "out")]),
}
METRIC = {Q4_0: "Q4_0 decode GEMV GB/s (Llama-3-8B shapes)", Q4_K: "Q4_K decode GEMV GB/s (Llama-3-70B shapes)",
          Q6_K: "Q6_K decode GEMV GB/s (Llama-3-8B shapes)"}
PREFILL_SHAPES = [(4096, 4096), (14336, 4096), (4096, 14336)]


def algorithmic_bytes(wtype, N, K, T=1):
    """N*K*bpe(weights) + T*K*4 (fp32 activations in) + T*N*4 (fp32 out)."""
    return N * K * BPE[wtype] + T * K * 4 + T * N * 4


def workload_config(name, layers=0):
    """The part of `config` both arms print identically."""
    wtype, L, layer, tail = WORKLOADS[name]
    L = layers or L
    mats = [(N, K) for _ in range(L) for (_, N, K, _) in layer] + [(N, K) for (_, N, K, _) in tail]
    return {"workload": name, "tokens": 1, "layers": L, "matmuls_per_step": len(mats),
            "weight_GB_per_step": sum(algorithmic_bytes(wtype, N, K) for N, K in mats) / 1e9,
            "l2": "distinct weights per step exceed L2 (126 MB) 30x; no flush needed"}


from llamafile_b200.shard import max_shard, shard_rows  # noqa: E402  (row split in 128-row tiles, as ggml-cuda.cu:646-674)


class ClockSampler(threading.Thread):
    """SM clock, power and throttle reasons of one GPU, sampled through NVML every ~2 ms (a decode step lasts ~1 ms: nvidia-smi's
    ~100 ms per query would never land inside the timed region). Falls back to polling nvidia-smi when pynvml is unusable."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag, self.marks = index, [], False, {}
        self.nv = None
        try:
            import pynvml
            pynvml.nvmlInit()
            # NVML enumerates physical devices: honour CUDA_VISIBLE_DEVICES when it is a plain index list
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            phys = int(vis.split(",")[index]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nv = pynvml
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:   # noqa: BLE001
            self.nv = None

    def run(self):
        if self.nv is not None:
            nv = self.nv
            R = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown, "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                 "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown, "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
            while not self.stop_flag:
                try:
                    sm = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                    mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                    pw = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
                    self.rows.append((sm, self.sm_max, pw, tuple(k for k, bit in R.items() if mask & bit)))
                except Exception:   # noqa: BLE001
                    pass
                time.sleep(0.002)
            return
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                c = [v.strip() for v in out.split(",")]
                self.rows.append((float(c[0]), float(c[1]), float(c[2]), tuple(n for n, v in zip(names, c[3:7]) if v.lower().startswith("active"))))
            except Exception:   # noqa: BLE001
                pass
            time.sleep(0.05)

    def mark(self, name):
        self.marks[name] = len(self.rows)

    def summary(self, lo=0, hi=None):
        rows = self.rows[lo:hi]
        reasons = sorted({r for row in rows for r in row[3]})
        return {"sm_mhz": statistics.median([r[0] for r in rows]) if rows else None, "sm_max_mhz": max(r[1] for r in rows) if rows else None,
                "reasons": reasons, "samples": len(rows), "power_w_max": max(r[2] for r in rows) if rows else None,
                "source": "nvml" if self.nv is not None else "nvidia-smi"}


def measured_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    out = {"hbm": (6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"), "tf_sustained": (1377.0, "fallback"), "tf_burst": (1630.9, "fallback")}
    if p.exists():
        try:
            j = json.loads(p.read_text())
            out["hbm"] = (float(j["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)")
            out["tf_sustained"] = (float(j["bf16_tflops_sustained"]), "measured (MEASURED_PEAKS.json bf16_tflops_sustained)")
            out["tf_burst"] = (float(j["bf16_tflops"]), "measured (MEASURED_PEAKS.json bf16_tflops)")
        except Exception:
            pass
    return out


# ------------------------------------------------------------------------------------------------ reference arm
class ReferenceStep:
    """The workload's FULL decode step (every matmul, every layer) through the reference's own CPU path: quantize_row_q8_0 /
    q8_K + llamafile_sgemm (iqk / tinyBLAS), all host threads, driven like llamafile/sgemm_matmul_test.cpp:32-40. One layer's
    weights are synthesized and copied to every layer's own host buffer (distinct memory, so nothing is cache-resident
    across layers — the same bytes the GPU arm streams)."""

    def __init__(self, workload, layers=0, threads=None):
        from oracle import binding as B
        self.B = B
        self.impl, self.kind = B.reference(), "reference"
        if self.impl is None:
            self.impl, self.kind = B.oracle(), "port"
        self.wtype, L, layer, tail = WORKLOADS[workload]
        self.layers = layers or L
        self.nth = threads or (os.cpu_count() or 1)
        rng = np.random.default_rng(0)

        def synth(N, K):
            ts, blck = TS_BLCK[self.wtype]
            nb = N * K // blck
            W = rng.integers(0, 256, (nb, ts), dtype=np.uint8)
            d = (rng.uniform(0.5, 1.5, nb) * 2e-3).astype(np.float16).view(np.uint8).reshape(-1, 2)
            if self.wtype == Q6_K:
                W[:, 208:210] = (rng.uniform(0.5, 1.5, nb) * 2e-5).astype(np.float16).view(np.uint8).reshape(-1, 2)
            else:
                W[:, :2] = d
            if self.wtype == Q4_K:
                W[:, 2:4] = (rng.uniform(0.5, 1.5, nb) * 4e-3).astype(np.float16).view(np.uint8).reshape(-1, 2)
            return W.reshape(-1)
        proto = [(N, K, synth(N, K)) for (_, N, K, _) in layer]
        self.mats = []
        for li in range(self.layers):
            for N, K, W in proto:
                self.mats.append((N, K, W if li == 0 else W.copy(), rng.standard_normal((1, K)).astype(np.float32)))
        for (_, N, K, _) in tail:
            self.mats.append((N, K, synth(N, K), rng.standard_normal((1, K)).astype(np.float32)))
        self.bytes = sum(algorithmic_bytes(self.wtype, N, K) for N, K, _, _ in self.mats)

    def step(self):
        for N, K, W, x in self.mats:
            if self.kind == "reference":
                self.impl.mul_mat(W, self.wtype, x, nth=self.nth)
            else:
                self.impl.mul_mat(W, self.wtype, x)

    def describe(self, reps):
        isa = f", ISA variant {self.impl.variant}" if self.kind == "reference" else ", OpenMP scalar port"
        return (f"full step: {self.layers} layers, {len(self.mats)} matmuls, {self.bytes / 1e9:.3f} GB of distinct host weights x {reps} reps, "
                f"activation quantization included{isa}")

    def timed(self, reps):
        t0 = time.perf_counter()
        for _ in range(reps):
            self.step()
        return (time.perf_counter() - t0) / reps


def run_reference_arm(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    t0 = time.perf_counter()
    ref = ReferenceStep(args.workload, args.layers)
    for _ in range(max(1, args.warmup)):
        ref.step()
    times = []
    for _ in range(args.steps):
        times.append(ref.timed(1))
        if time.perf_counter() - t0 > 150 and len(times) >= 3:   # bounded: the run ends within a few minutes on any host
            break
    dt = statistics.median(times)
    value = ref.bytes / dt / 1e9
    line = {"impl": "reference", "metric": METRIC[ref.wtype], "value": value, "unit": "GB/s", "n_gpus": args.gpus, "steps": len(times),
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "int8 block dot products, fp32 accumulate", "data": "synthetic", "config": workload_config(args.workload, args.layers),
            "cpu_baseline": {"value": value, "unit": "GB/s", "cores": ref.nth if ref.kind == "reference" else (os.cpu_count() or 1),
                             "kind": ref.kind, "sample": ref.describe(len(times))},
            "e2e": {"value": value, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0,
            "decode_tokens_per_s": 1.0 / dt, "wall_s": time.perf_counter() - t0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ B200 arm
def log(msg):
    if os.environ.get("TB200_BENCH_VERBOSE"):
        sys.stderr.write(f"[bench r{os.environ.get('RANK', '0')} {time.strftime('%H:%M:%S')}] {msg}\n")
        sys.stderr.flush()


def synth_dev(torch, dev, wtype, N, K, seed):
    """Random quant bytes with realistic fp16 scales, built on the device."""
    g = torch.Generator(device=dev); g.manual_seed(seed)
    if wtype == F16:
        return (torch.randn(N, K, device=dev, generator=g) * 0.02).to(torch.float16).view(torch.uint8).reshape(-1)
    ts, blck = TS_BLCK[wtype]
    nb = N * K // blck
    raw = torch.randint(0, 256, (nb, ts), dtype=torch.uint8, device=dev, generator=g)
    d = (torch.rand(nb, device=dev, generator=g) + 0.5) * 2e-3
    raw[:, 0:2] = d.to(torch.float16).view(torch.uint8).reshape(nb, 2)
    if wtype == Q4_K:
        raw[:, 2:4] = (d * 2).to(torch.float16).view(torch.uint8).reshape(nb, 2)
    if wtype == Q6_K:   # block_q6_K keeps d in its last two bytes; the first two are quants
        raw[:, 0:2] = torch.randint(0, 256, (nb, 2), dtype=torch.uint8, device=dev, generator=g)
        raw[:, 208:210] = (d * 1e-2).to(torch.float16).view(torch.uint8).reshape(nb, 2)
    return raw.reshape(-1)


def sustained(torch, stream, fn, min_s=1.0, calib=3):
    """fn() launched back to back for >= min_s (the sustained regime: clocks and power settle); ms per call by CUDA events."""
    with torch.cuda.stream(stream):
        for _ in range(calib):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(calib):
            fn()
        e1.record(stream)
        stream.synchronize()
        per = max(e0.elapsed_time(e1) / calib, 1e-3)
        n = max(5, int(min_s * 1e3 / per) + 1)
        e0.record(stream)
        for _ in range(n):
            fn()
        e1.record(stream)
        stream.synchronize()
    return e0.elapsed_time(e1) / n, n


def prefill_leg(torch, tb, dev, stream, sampler, peaks):
    """BASELINE config 3: Q8_0 and Q4_0 prefill GEMM on tcgen05, T in {512, 4096}, the three Llama-3-8B shapes. Each point runs
    tb200_mul_mat (activation quantize + fp16 prep + GEMM) back to back for >= 1 s; weights/outputs rotate over enough
    distinct copies to exceed L2 where one problem would fit in it."""
    out = []
    for wtype, name in ((Q8_0, "q8_0"), (Q4_0, "q4_0")):
        for (N, K) in PREFILL_SHAPES:
            for T in (512, 4096):
                foot = N * K * BPE[wtype] + T * K * 4 + T * N * 4
                copies = max(1, min(8, int(300e6 // foot) + 1))
                ws = [tb.Weights(synth_dev(torch, dev, wtype, N, K, seed=c), wtype, N, K) for c in range(copies)]
                x = torch.randn(T, K, device=dev)
                outs = [torch.empty(T, N, device=dev) for _ in range(copies)]
                it = [0]

                def fn():
                    c = it[0] % copies
                    it[0] += 1
                    tb.mul_mat(ws[c], x, T, outs[c])
                sampler.mark(f"pf{len(out)}a")
                ms, n = sustained(torch, stream, fn, min_s=1.0)
                sampler.mark(f"pf{len(out)}b")
                tf = 2.0 * T * N * K / (ms * 1e-3) / 1e12
                clk = sampler.summary(sampler.marks[f"pf{len(out)}a"], sampler.marks[f"pf{len(out)}b"])
                out.append({"type": name, "N": N, "K": K, "T": T, "ms": ms, "calls": n, "TFLOPS": tf, "frac_of_measured_sustained": tf / peaks["tf_sustained"][0],
                            "frac_of_spec_2250": tf / SPEC_BF16_TFLOPS, "distinct_copies": copies, "sm_mhz": clk["sm_mhz"], "power_w_max": clk["power_w_max"],
                            "reasons": clk["reasons"]})
                log(f"prefill {name} N={N} K={K} T={T}: {tf:.0f} TFLOPS")
                for w in ws:
                    w.free()
                del outs, x
    return out


def mixmul_leg(torch, tb, dev, stream):
    """BASELINE config 4: Mixtral-8x7B ffn_gate_exps (8 experts x 14336 x 4096, Q4_0), top-2, through tb200_mixmul_dev
    (activation quantize + device-side expert grouping + grouped kernel); 264 MB of experts exceed L2."""
    rows, cols, experts, top = 14336, 4096, 8, 2
    w = tb.Weights(synth_dev(torch, dev, Q4_0, rows * experts, cols, seed=5), Q4_0, rows * experts, cols)
    g = torch.Generator(device=dev); g.manual_seed(1)
    res = []
    for tokens in (1, 512):
        x = torch.randn(tokens, 1, cols, device=dev, generator=g)
        plan = torch.stack([torch.randperm(experts, device=dev, generator=g)[:top] for _ in range(tokens)]).to(torch.int32).contiguous()
        outb = torch.empty(tokens, top, rows, device=dev)
        touched = int(torch.unique(plan).numel())
        ms, n = sustained(torch, stream, lambda: tb.mixmul(w, experts, x, 1, plan, top, tokens, outb), min_s=0.5)
        byts = touched * rows * cols * BPE[Q4_0] + tokens * cols * 4 + tokens * top * rows * 4
        res.append({"op": "mul_mat_id", "type": "q4_0", "rows": rows, "cols": cols, "experts": experts, "top": top, "tokens": tokens,
                    "experts_touched": touched, "ms": ms, "calls": n, "GBps": byts / (ms * 1e-3) / 1e9, "TFLOPS": 2.0 * tokens * top * rows * cols / (ms * 1e-3) / 1e12})
    w.free()
    return res


def quantizer_leg(torch, tb, dev, stream, peak_hbm):
    """BASELINE.md §3 last row: T x K fp32 -> Q8_0 blocks, T=512, K=4096, rotating over 32 distinct inputs (268 MB > L2)."""
    T, K, copies = 512, 4096, 32
    xs = [torch.randn(T, K, device=dev) for _ in range(copies)]
    y = torch.empty(T * K // 32 * 34, dtype=torch.uint8, device=dev)
    it = [0]

    def fn():
        tb.quantize_q8_0(xs[it[0] % copies], T, K, y)
        it[0] += 1
    ms, n = sustained(torch, stream, fn, min_s=0.3)
    byts = T * K * 4 + T * K // 32 * 34
    return {"op": "quantize_row_q8_0", "rows": T, "k": K, "ms": ms, "calls": n, "GBps": byts / (ms * 1e-3) / 1e9, "frac": byts / (ms * 1e-3) / 1e9 / peak_hbm}


def run_b200_arm(args):
    import torch
    import torch.distributed as dist

    log("start")
    # stdout carries ONE JSON line: native libraries (NCCL's banner, whatever the caller's NCCL_DEBUG asks for) write to fd 1
    # directly, so fd 1 is pointed at stderr for the whole run and the line goes out through a saved copy of the real stdout
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from llamafile_b200.build import build
    build()
    import llamafile_b200 as tb
    if not tb.link():
        raise SystemExit("bench: no sm_100 device; the CUDA path cannot run and there is no CPU fallback")
    log("library linked")
    dev = torch.device("cuda", local)
    stream = torch.cuda.Stream(device=dev)
    tb.set_stream(stream.cuda_stream)
    peaks = measured_peaks()

    wtype, layers, layer, tail = WORKLOADS[args.workload]
    if args.layers:
        layers = args.layers
    use_plan = args.decode == "plan" and wtype in (Q4_0, Q4_K)   # N > 1: the replicated groups run as plan launches, the sharded ones through the fused all-gather

    specs = [(f"l{li}.{n}", N, K, f"l{li}.{g}") for li in range(layers) for (n, N, K, g) in layer] + [(n, N, K, g) for (n, N, K, g) in tail]
    # north_star: "weight matrices of layers large enough to split are sharded by output rows". A launch group (the matrices
    # sharing one input) is split when its weights exceed --split-min-mb; smaller groups are computed in full on every rank.
    if args.split_min_mb is None:
        # Break-even of a split: it saves bytes*(1 - 1/W) / 6.5 TB/s of streaming and costs one cross-GPU exchange on the group's
        # critical path - ~3.2..4 us inside the multi-GPU plan (remote exchange round trip + flag + poll), ~6 us as
        # kernel -> barrier kernel -> kernel on the one-shot path (round 1's 80 MB rule). On 2 GPUs a 66 MB group (gate + up) measured
        # as a net loss (1.040 vs 0.993 ms per 8B step) although the formula says +2 us: its halves are chain-bound, not
        # bandwidth-bound, so W = 2 only splits what is far above the break-even (the vocabulary matrix).
        in_plan = use_plan and args.gather == "p2p"
        cost_us = ({4: 3.6}.get(world, 4.0) if in_plan else 6.0)
        args.split_min_mb = round(cost_us * 6.5 / (1.0 - 1.0 / world), 1) if world > 1 else 80.0   # us x 6.5 TB/s = MB
        if in_plan and world == 2 and wtype == Q4_0:
            # measured on 2 GPUs (8B step): everything replicated with the single-GPU kernel 0.971 ms; only the vocabulary matrix
            # split 0.999 ms; gate/up split as well 1.040 ms. Two ranks do not pay for an exchange at T = 1: nothing is split.
            args.split_min_mb = 1e6
        if not in_plan:
            args.split_min_mb = 80.0
    grp_bytes = {}
    for (name, N, K, grp) in specs:
        grp_bytes[grp] = grp_bytes.get(grp, 0) + N * K * BPE[wtype]
    split_grp = {g: world > 1 and b >= args.split_min_mb * 1e6 for g, b in grp_bytes.items()}
    mats, host_raw = [], []   # (name, N_full, lo, hi, K, group, Weights)

    def shard_seed(i, r, grp):
        return 1000 * i + (r if split_grp[grp] else 0)   # replicated groups hold the same weights on every rank
    with torch.cuda.stream(stream):
        for i, (name, N, K, grp) in enumerate(specs):
            lo, hi = shard_rows(N, rank, world) if split_grp[grp] else (0, N)
            raw = synth_dev(torch, dev, wtype, hi - lo, K, seed=shard_seed(i, rank, grp))
            w = tb.Weights(raw, wtype, hi - lo, K)
            if args.e2e:
                host_raw.append(raw.cpu().numpy())
            del raw
            mats.append((name, N, lo, hi, K, grp, w))
    torch.cuda.synchronize()

    # one fp32 input per launch group (attn / o / ffn / down per layer, + output); matrices that consume the same input
    # go out together (tb200_mul_mat_acts_multi / one plan group: wq/wk/wv, gate/up)
    groups, schedule = {}, []
    for i, m in enumerate(mats):
        if m[5] not in groups:
            gx = torch.Generator(device=dev); gx.manual_seed(77000 + len(groups))   # the activations are replicated: same values on every rank
            groups[m[5]] = (torch.randn(1, m[4], device=dev, dtype=torch.float32, generator=gx),
                            None if (use_plan and not split_grp[m[5]]) else tb.Acts(wtype, m[4], 1))
        if schedule and schedule[-1][0] == m[5] and len(schedule[-1][1]) < 4 and wtype in (Q4_0, Q4_K, Q6_K):
            schedule[-1][1].append(i)
        else:
            schedule.append((m[5], [i]))
    # outputs: each rank writes its row slices of a launch group into ONE staging row (every slice padded to the widest shard)
    outs, staging, gathered = [None] * len(mats), [], []
    for grp, idx in schedule:
        widths = [max_shard(mats[i][1], world) if split_grp[grp] else mats[i][1] for i in idx]
        st = torch.zeros(1, sum(widths), device=dev, dtype=torch.float32)
        off = 0
        for i, wd in zip(idx, widths):
            outs[i] = st[:, off: off + (mats[i][3] - mats[i][2])]
            off += wd
        staging.append(st)
        gathered.append(torch.empty(world, st.shape[1], device=dev, dtype=torch.float32) if world > 1 and args.gather == "nccl" else None)

    pg, p2p_off = None, None
    if world > 1 and args.gather == "p2p":
        from llamafile_b200.peer import PeerGroup
        p2p_off, off = [], 0
        for (_, N, lo, hi, K, g, w) in mats:
            p2p_off.append(off)
            off += N * 4
        pg = PeerGroup(world, rank, off)

    plan, plan_index, segments = None, {}, []
    sharded_plan = (use_plan and world > 1 and pg is not None and os.environ.get("TB200_BENCH_SHARDED_PLAN", "1") != "0" and any(split_grp.values()) and
                    all(mats[i][3] > mats[i][2] for grp, idx in schedule if split_grp[grp] for i in idx))
    if sharded_plan:
        # N > 1: the whole step as ONE multi-GPU plan (tb200_plan_create_peers). A split group's outputs point at this rank's row
        # slice inside the symmetric buffer: the kernel stores them into every rank's copy and the barrier after the group spans
        # the GPUs; replicated groups keep private outputs and a local barrier.
        plan = tb.Plan([{"w": [mats[i][6] for i in idx],
                         "out": [(pg.out + p2p_off[i] + mats[i][2] * 4) if split_grp[grp] else outs[i] for i in idx],
                         "x": groups[grp][0]} for grp, idx in schedule], peers=pg)
        plan_index = {gi: gi for gi in range(len(schedule))}
        segments = [["plan", 0, len(schedule)]]
    elif use_plan:
        planned = [gi for gi, (grp, idx) in enumerate(schedule) if not split_grp[grp]]
        plan = tb.Plan([{"w": [mats[i][6] for i in schedule[gi][1]], "out": [outs[i] for i in schedule[gi][1]], "x": groups[schedule[gi][0]][0]} for gi in planned])
        plan_index = {gi: pi for pi, gi in enumerate(planned)}
        # the step as segments: maximal runs of consecutive replicated groups (one plan launch each) and single sharded groups
        for gi in range(len(schedule)):
            if gi in plan_index and segments and segments[-1][0] == "plan" and segments[-1][2] == gi:
                segments[-1][2] = gi + 1
            elif gi in plan_index:
                segments.append(["plan", gi, gi + 1])
            else:
                segments.append(["split", gi, gi + 1])

    def step(with_gather=True, with_quant=True):
        if plan is not None and len(segments) == 1:
            plan.run()
            return
        todo = segments if plan is not None else [["split", 0, len(schedule)]]
        for kind, g0, g1 in todo:
            if kind == "plan":
                plan.run(plan_index[g0], g1 - g0)
                continue
            run_groups(range(g0, g1), with_gather, with_quant)

    def run_groups(gis, with_gather, with_quant):
        for gi in gis:
            grp, idx = schedule[gi]
            x, acts = groups[grp]
            if with_quant:
                acts.quantize(x, 1)
            if pg is not None and with_gather and split_grp[grp]:
                live = [i for i in idx if mats[i][3] > mats[i][2]]
                if live:
                    pg.mul_mat_gather([mats[i][6] for i in live], acts, [p2p_off[i] for i in live], [mats[i][1] for i in live],
                                      [mats[i][2] for i in live])
                pg.barrier()
                continue
            tb.mul_mat_acts_multi([mats[i][6] for i in idx], acts, [outs[i] for i in idx])
            if world > 1 and with_gather and split_grp[grp]:
                dist.all_gather_into_tensor(gathered[gi], staging[gi])

    step_bytes = sum(algorithmic_bytes(wtype, hi - lo, K) for (_, N, lo, hi, K, g, w) in mats)
    total_step_bytes = sum(algorithmic_bytes(wtype, N, K) for (_, N, lo, hi, K, g, w) in mats)
    graphs = []

    def capture(fn):
        """CUDA graph of one step on our stream (falls back to eager launches if capture is refused)."""
        with torch.cuda.stream(stream):
            fn()   # grows nothing afterwards; sets kernel attributes
            stream.synchronize()
            if args.no_graph:
                return fn, False
            try:
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=stream):
                    fn()
                graphs.append(g)
                return g.replay, True
            except Exception as e:   # noqa: BLE001
                sys.stderr.write(f"bench: graph capture unavailable ({e}); timing eager launches\n")
                torch.cuda.synchronize()
                return fn, False

    def timed(run, steps, warmup):
        with torch.cuda.stream(stream):
            for _ in range(warmup):
                run()
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(steps):
                run()
            e1.record(stream)
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    log('weights resident; first step')
    l0 = tb.launch_count()
    with torch.cuda.stream(stream):
        step()
    stream.synchronize()
    launches_per_step = tb.launch_count() - l0
    if plan is not None and plan.status() != 0:
        raise SystemExit("bench: the decode plan kernel gave up waiting (status != 0)")

    # ---- N > 1: the exchanged outputs must be the unsharded product. Rank 0 regenerates every rank's shard of a few split
    # matrices from their seeds, multiplies them unsharded on its own GPU and compares sampled rows with its copy of the
    # symmetric buffer (fused path) / the gathered rows (NCCL path).
    parity_checked = None
    if world > 1:
        ok, checked = True, 0
        split_idx = [i for i, m in enumerate(mats) if split_grp[m[5]]]
        for i in ([split_idx[0], split_idx[len(split_idx) // 2], split_idx[-1]] if split_idx else []):
            name, N, lo, hi, K, grp, w = mats[i]
            if rank == 0:
                x = groups[grp][0]
                parts = []
                for r in range(world):
                    rlo, rhi = shard_rows(N, r, world)
                    if rhi > rlo:
                        with torch.cuda.stream(stream):   # the module's launches go to `stream`: generate the bytes there too (the 300 MB vocabulary shard was repacked before torch's default stream had written it)
                            raw = synth_dev(torch, dev, wtype, rhi - rlo, K, seed=shard_seed(i, r, grp))
                            wr = tb.Weights(raw, wtype, rhi - rlo, K)
                            o = torch.empty(1, rhi - rlo, device=dev)
                            tb.mul_mat(wr, x, 1, o)
                        stream.synchronize()
                        parts.append(o.cpu().numpy()[0])
                        wr.free()
                full = np.concatenate(parts)
                if pg is not None:
                    got = pg.read(p2p_off[i], N)
                else:
                    gi = next(k for k, (_, idx) in enumerate(schedule) if i in idx)
                    col = sum(max_shard(mats[j][1], world) for j in schedule[gi][1] if j < i)
                    gg = gathered[gi].cpu().numpy()
                    got = np.concatenate([gg[r, col: col + (shard_rows(N, r, world)[1] - shard_rows(N, r, world)[0])] for r in range(world)])
                den = float(np.abs(full).max()) or 1.0
                good = (not np.isnan(got).any()) and float(np.abs(got - full).max()) / den <= 1e-3
                ok, checked = ok and good, checked + 1
                if not good:
                    bad = np.nonzero(~(np.abs(got - full) <= 1e-3 * den))[0]
                    sys.stderr.write(f"bench: PARITY MISMATCH on gathered output of {name}: {bad.size} of {N} rows differ, first {bad[:8].tolist()}, last {bad[-4:].tolist()}, "
                                     f"nan {int(np.isnan(got).sum())}, got {got[bad[:4]].tolist()} want {full[bad[:4]].tolist()}\n")
        flag = torch.tensor([1 if ok else 0], device=dev)
        dist.broadcast(flag, 0)
        if int(flag.item()) != 1 and os.environ.get("TB200_BENCH_IGNORE_PARITY") != "1":   # (the override is for kernel experiments that break results on purpose)
            raise SystemExit("bench: gathered outputs differ from the unsharded product")
        parity_checked = None if int(flag.item()) != 1 else {"matrices": 3 if split_idx else 0, "rows": "all", "tolerance": 1e-3, "against": "the same matrices multiplied unsharded on rank 0"}

    log('capturing full step')
    run_full, graphed = capture(step)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    log('timing full step')
    sampler.mark("step_a")
    ms_total = timed(run_full, args.steps, args.warmup)
    ms_per_step = ms_total / args.steps
    value = total_step_bytes / (ms_per_step * 1e-3) / 1e9

    log('roofline pass')
    peak, peak_src = peaks["hbm"]
    if plan is not None and len(segments) == 1:
        # the plan kernel IS the step: one launch, quantization inside
        ms_kernel, n_launch, kname = ms_per_step, 1, "plan_kernel<%s>" % {Q4_0: "Q4_0", Q4_K: "Q4_K"}[wtype]
    elif plan is not None:
        # N > 1: the plan launches of the replicated groups + the sharded groups' GEMVs, no exchange and no quantizer launches
        run_gemv, _ = capture(lambda: step(with_gather=False, with_quant=False))
        ms_kernel = timed(run_gemv, args.steps, args.warmup) / args.steps
        n_launch, kname = len(segments), "plan_kernel + gemv kernel (sharded groups)"
    else:
        run_gemv, _ = capture(lambda: step(with_gather=False, with_quant=False))
        ms_kernel = timed(run_gemv, args.steps, args.warmup) / args.steps
        n_launch, kname = len(schedule), {Q4_0: "gemv_q4_0_kernel", Q4_K: "gemv_q4_K_kernel", Q6_K: "gemv_q6_K_kernel"}[wtype]
    sampler.mark("step_b")
    achieved = step_bytes / (ms_kernel * 1e-3) / 1e9
    # DRAM traffic per launch: (dram read + write) / algorithmic bytes of the same kernel from the committed ncu --set full capture,
    # applied to this run's algorithmic bytes — derived, not measured in this run
    traffic, traffic_src = None, None
    try:
        tj = json.loads((ROOT / "profiles" / ("r02_plan_traffic.json" if plan is not None else "r01_gemv_traffic.json")).read_text())
        if wtype == Q4_0:
            traffic, traffic_src = tj["traffic_over_algorithmic"] * step_bytes / n_launch, tj["source"]
    except Exception:   # noqa: BLE001
        pass
    roofline = {"bound": "hbm", "kernel": kname, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src,
                "traffic": traffic, "traffic_derived": True, "traffic_source": traffic_src, "launches_timed": n_launch,
                "avg_launch_us": ms_kernel * 1e3 / n_launch, "algorithmic_bytes_per_launch": step_bytes / n_launch,
                "frac_of_spec_8TBs": achieved / SPEC_HBM_GBS}
    log('per-shape pass')
    # per-shape view (BASELINE configs[1] names N=4096/14336/128256 at K=4096): every matrix of one shape, each its own dependent group
    per_shape = {}
    for shape in sorted({(hi - lo, K) for (_, N, lo, hi, K, g, w) in mats}):
        idx = [i for i, m in enumerate(mats) if (m[3] - m[2], m[4]) == shape]
        sub = None
        if plan is not None and not split_grp[mats[idx[0]][5]]:
            sub = tb.Plan([{"w": [mats[i][6]], "out": [outs[i]], "x": groups[mats[i][5]][0]} for i in idx])
            run_s = sub.run
        else:
            def only(idx=idx):
                for i in idx:
                    tb.mul_mat_acts(mats[i][6], groups[mats[i][5]][1], outs[i])
            run_s, _ = capture(only)
        ms = timed(run_s, max(3, args.steps // 2), 2) / max(3, args.steps // 2)
        b = algorithmic_bytes(wtype, shape[0], shape[1]) * len(idx)
        per_shape[f"N={shape[0]},K={shape[1]}"] = {"matrices": len(idx), "us_per_matrix": ms * 1e3 / len(idx), "GBps": b / (ms * 1e-3) / 1e9,
                                                  "frac": b / (ms * 1e-3) / 1e9 / peak, "distinct_MB": b / 1e6}
        if sub is not None:
            sub.free()
            # the same matrices as INDEPENDENT groups (TB200_PLAN_INDEPENDENT: every matrix has its own weights, input and output,
            # nothing depends on anything): the GEMV path's streaming rate for this shape without a dependency chain between matrices
            sub = tb.Plan([{"w": [mats[i][6]], "out": [outs[i]], "x": groups[mats[i][5]][0], "prologue": tb.PLAN_QUANT | tb.PLAN_INDEPENDENT} for i in idx])
            ms_i = timed(sub.run, max(3, args.steps // 2), 2) / max(3, args.steps // 2)
            per_shape[f"N={shape[0]},K={shape[1]}"].update({"independent_us_per_matrix": ms_i * 1e3 / len(idx), "independent_GBps": b / (ms_i * 1e-3) / 1e9,
                                                            "independent_frac": b / (ms_i * 1e-3) / 1e9 / peak})
            sub.free()
    roofline["per_shape"] = per_shape
    roofline["per_shape_note"] = ("us_per_matrix / GBps / frac: every matrix of the shape as its own group of ONE launch, each group waiting for the one "
                                  "before it (a decode step's dependency chain); independent_*: the same launch with the groups declared independent")

    extra_clock_marks = {}
    if world == 1 and rank == 0 and not args.no_extra:
        log('prefill leg')
        roofline["prefill"] = prefill_leg(torch, tb, dev, stream, sampler, peaks)
        roofline["prefill_peak"] = {"bf16_tflops_sustained": peaks["tf_sustained"][0], "bf16_tflops_burst": peaks["tf_burst"][0], "spec": SPEC_BF16_TFLOPS,
                                    "source": peaks["tf_sustained"][1], "includes": "activation quantize + fp16 prep + tcgen05 GEMM, back to back >= 1 s per point"}
        log('mixmul leg')
        roofline["mixmul"] = mixmul_leg(torch, tb, dev, stream)
        roofline["quantizer"] = quantizer_leg(torch, tb, dev, stream, peak)

    log('e2e pass')
    # ---- e2e (1): the module as a ggml backend (include/tb200_ggml_backend.h; what llamafile/cuda.c would load). The step's 225
    # matmuls are ONE ggml graph of dependent MUL_MAT nodes (x -> wq/wk/wv -> wo -> gate/up -> down -> next layer ... -> output);
    # per step the fp32 hidden state goes host -> device from pinned memory (set_tensor_async), graph_compute runs the graph
    # without returning to the host (consecutive one-token Q4_0 nodes = one persistent kernel), the logits come back
    # (get_tensor_async) and the host synchronizes. Weights were uploaded once with set_tensor, as a model load does.
    e2e = None
    if args.e2e and world == 1:
        from llamafile_b200.ggml_backend import StandaloneHost
        ts, blck = TS_BLCK[wtype]
        raw_bytes = sum(N * K // blck * ts + 512 for (_, N, K, _) in specs)
        host = StandaloneHost(local, raw_bytes + (64 << 20))

        def chain_weights(N, K, seed):
            """Synthetic weights under which a hidden state keeps its magnitude through 129 dependent matmuls: zero-mean quants
            (Q4_0 nibbles 1..15) and a block scale of 1/sqrt(var(q) * K); for the other types the generic synthetic bytes."""
            if wtype != Q4_0:
                return synth_dev(torch, dev, wtype, N, K, seed)
            g = torch.Generator(device=dev); g.manual_seed(seed)
            nb = N * K // 32
            raw = torch.empty((nb, 18), dtype=torch.uint8, device=dev)
            lo = torch.randint(1, 16, (nb, 16), dtype=torch.uint8, device=dev, generator=g)
            hi = torch.randint(1, 16, (nb, 16), dtype=torch.uint8, device=dev, generator=g)
            raw[:, 2:] = lo | (hi << 4)
            d = (torch.rand(nb, device=dev, generator=g) * 0.2 + 0.9) / (18.67 * K) ** 0.5
            raw[:, 0:2] = d.to(torch.float16).view(torch.uint8).reshape(nb, 2)
            return raw.reshape(-1)
        wt_tensors = {}
        with torch.cuda.stream(stream):
            for i, (name, N, K, grp) in enumerate(specs):
                t = host.tensor(wtype, K, N)
                raw = chain_weights(N, K, 5000 + i)
                stream.synchronize()
                host.set(t, raw.data_ptr(), raw.numel())
                wt_tensors[name] = t
                del raw
        K0 = specs[0][2]
        x_t = host.tensor(F32, K0, 1)
        nodes, cur = [], x_t
        for li in range(layers):
            W = lambda n: wt_tensors[f"l{li}.{n}"]   # noqa: E731
            q, k_, v = host.mul_mat(W("wq"), cur), host.mul_mat(W("wk"), cur), host.mul_mat(W("wv"), cur)
            o = host.mul_mat(W("wo"), q)
            ga, up = host.mul_mat(W("gate"), o), host.mul_mat(W("up"), o)
            cur = host.mul_mat(W("down"), ga)
            nodes += [q, k_, v, o, ga, up, cur]
        logits_t = host.mul_mat(wt_tensors["output"], cur)
        nodes.append(logits_t)
        graph = host.graph(nodes)
        n_logits = specs[-1][1]
        hx = torch.randn(K0, dtype=torch.float32).pin_memory()
        hl = torch.zeros(n_logits, dtype=torch.float32).pin_memory()

        def backend_step():
            host.set_async(x_t, hx.data_ptr(), K0 * 4)
            if host.compute(graph) != 0:
                raise SystemExit("bench: graph_compute failed")
            host.get_async(logits_t, hl.data_ptr(), n_logits * 4)
            host.synchronize()
        for _ in range(max(3, args.warmup)):
            backend_step()
        be_launches, be_plan_nodes = host.stats()
        lg = hl.numpy()
        if not (np.isfinite(lg).all() and float(np.abs(lg).max()) > 0):
            raise SystemExit("bench: backend e2e produced degenerate logits")
        steps_e = max(5, args.steps)
        t0 = time.perf_counter()
        for _ in range(steps_e):
            backend_step()
        dt = (time.perf_counter() - t0) / steps_e
        e2e = {"value": total_step_bytes / dt / 1e9, "unit": "GB/s", "h2d_bytes_per_step": K0 * 4, "d2h_bytes_per_step": n_logits * 4,
               "ms_per_step": dt * 1e3, "kernel_launches_per_step": be_launches, "nodes_in_decode_plan": be_plan_nodes, "graph_nodes": len(nodes),
               "logits_rms": float(np.sqrt(np.mean(lg.astype(np.float64) ** 2))),
               "api": "ggml backend object (tb200_backend_init): set_tensor_async(hidden state, pinned host) -> graph_compute(225 dependent MUL_MAT nodes) -> "
                      "get_tensor_async(logits, pinned host) -> synchronize; weights uploaded once with set_tensor"}
        # the same weights as a llama-shaped graph WITH the matmuls' neighbours (llm_build_norm / llm_build_ffn, no attention):
        # per layer h -> rms_norm*w -> q,k,v ; o = wo q + h ; rms_norm*w -> gate,up ; silu(gate)*up -> down + o ; then norm -> output.
        # The backend folds every neighbour into the plan: still one kernel launch per step, no fp32 intermediate in memory.
        try:
            ones = torch.ones(max(K0, 1), dtype=torch.float32, device=dev)
            nw_t = host.tensor(F32, K0, 1)
            host.set(nw_t, ones.data_ptr(), K0 * 4)
            nodes2, h = [], x_t
            for li in range(layers):
                W = lambda n: wt_tensors[f"l{li}.{n}"]   # noqa: E731
                n1 = host.rms_norm(h, 1e-5); c1 = host.mul(n1, nw_t)
                q, k_, v = host.mul_mat(W("wq"), c1), host.mul_mat(W("wk"), c1), host.mul_mat(W("wv"), c1)
                o = host.mul_mat(W("wo"), q); h2 = host.add(o, h)
                n2 = host.rms_norm(h2, 1e-5); c2 = host.mul(n2, nw_t)
                ga, up = host.mul_mat(W("gate"), c2), host.mul_mat(W("up"), c2)
                si = host.silu(ga); act = host.mul(si, up)
                d = host.mul_mat(W("down"), act); h = host.add(d, h2)
                nodes2 += [n1, c1, q, k_, v, o, h2, n2, c2, ga, up, si, act, d, h]
            nf = host.rms_norm(h, 1e-5); cf = host.mul(nf, nw_t)
            logits2 = host.mul_mat(wt_tensors["output"], cf)
            nodes2 += [nf, cf, logits2]
            graph2 = host.graph(nodes2)

            def block_step():
                host.set_async(x_t, hx.data_ptr(), K0 * 4)
                if host.compute(graph2) != 0:
                    raise SystemExit("bench: graph_compute (block graph) failed")
                host.get_async(logits2, hl.data_ptr(), n_logits * 4)
                host.synchronize()
            for _ in range(max(3, args.warmup)):
                block_step()
            l2, p2 = host.stats()
            ok2 = bool(np.isfinite(hl.numpy()).all())
            t0 = time.perf_counter()
            for _ in range(steps_e):
                block_step()
            dt2 = (time.perf_counter() - t0) / steps_e
            e2e["with_neighbours"] = {"ms_per_step": dt2 * 1e3, "value": total_step_bytes / dt2 / 1e9, "unit": "GB/s", "graph_nodes": len(nodes2),
                                      "kernel_launches_per_step": l2, "matmuls_in_decode_plan": p2, "finite": ok2,
                                      "graph": "per layer: rms_norm, mul(norm weight), wq/wk/wv, wo, add(residual), rms_norm, mul, gate/up, silu, mul, down, add; final norm + output matrix"}
        except Exception as ex:   # noqa: BLE001  (an extra, never the headline)
            e2e["with_neighbours"] = {"error": str(ex)[:200]}
        host.free()
        del host

    # ---- e2e (2): llamafile_sgemm() with HOST buffers, one blocking call per matmul (the thread-0 shim of INTEGRATION.md §1);
    # weights registered as the immutable mmap'd range and resident after first use
    if args.e2e:
        vdt, vts = (Q8_0, 34) if wtype == Q4_0 else (15, 292)   # Q8_K activations for the k-quants
        blck = TS_BLCK[wtype][1]
        hB, hC = [], []
        for i, (name, N, lo, hi, K, grp, w) in enumerate(mats):
            x = groups[grp][0].cpu().numpy()
            q = tb.quantize_row_q8_0(x[0]) if wtype == Q4_0 else tb.quantize_row_q8_K(x[0])
            b = torch.from_numpy(q)
            c = torch.empty(hi - lo, dtype=torch.float32)
            if args.e2e_pinned:
                b, c = b.pin_memory(), c.pin_memory()
            hB.append(b); hC.append(c)
        for raw in host_raw:
            tb.register_weights(raw)
        kb = lambda K: K // blck   # noqa: E731
        from llamafile_b200.lib import lib as _lib
        sgemm = _lib().llamafile_sgemm
        calls = [(hi - lo, 1, kb(K), host_raw[i].ctypes.data, kb(K), hB[i].data_ptr(), kb(K), hC[i].data_ptr(), hi - lo, 0, 1, wtype, vdt, F32)
                 for i, (name, N, lo, hi, K, grp, w) in enumerate(mats)]

        def host_step():
            for a in calls:
                if not sgemm(*a):
                    raise SystemExit("bench: llamafile_sgemm returned false")
        for _ in range(max(2, args.warmup)):
            host_step()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        steps_e = max(2, args.steps // 2)
        t0 = time.perf_counter()
        for _ in range(steps_e):
            host_step()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / steps_e
        if world > 1:
            t = torch.tensor([dt], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX); dt = float(t.item())
        shim = {"value": total_step_bytes / dt / 1e9, "unit": "GB/s", "h2d_bytes_per_step": int(sum(kb(m[4]) * vts for m in mats)),
                "d2h_bytes_per_step": int(sum((m[3] - m[2]) * 4 for m in mats)), "ms_per_step": dt * 1e3,
                "api": "llamafile_sgemm(host A registered as immutable and resident on the device, host B Q8_0, host C; "
                       + ("pinned" if args.e2e_pinned else "pageable, as stock ggml buffers are") + "), one blocking call per matmul"}
        for raw in host_raw:
            tb.unregister_weights(raw)
        if e2e is None:
            e2e = shim
        else:
            e2e["llamafile_sgemm_shim"] = shim

    log('done; assembling line')
    clocks = None
    if rank == 0:
        sampler.stop_flag = True
        sampler.join(timeout=2)
        clocks = sampler.summary(sampler.marks.get("step_a", 0), sampler.marks.get("step_b"))
        clocks["whole_run"] = sampler.summary()

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            ref = ReferenceStep(args.workload, args.layers)
            ref.step()
            reps, t0 = 0, time.perf_counter()
            while reps < 2 or time.perf_counter() - t0 < 10.0:
                ref.step()
                reps += 1
            dt = (time.perf_counter() - t0) / reps
            cpu_baseline = {"value": ref.bytes / dt / 1e9, "unit": "GB/s", "cores": ref.nth if ref.kind == "reference" else (os.cpu_count() or 1),
                            "kind": ref.kind, "sample": ref.describe(reps)}
            del ref
        except Exception as e:   # noqa: BLE001
            cpu_baseline = {"value": None, "unit": "GB/s", "cores": 0, "kind": "port", "sample": f"failed: {e}"}

    timed_out = pg is not None and pg.timed_out()
    if plan is not None and plan.status() != 0:
        timed_out = True
    if rank == 0 and not timed_out:
        cfg = workload_config(args.workload, args.layers)
        cfg.update({"launch_groups_per_step": len(schedule), "launches_per_step": int(launches_per_step),
                    "decode": (("multi-GPU plan: one persistent kernel per step and rank; split groups store their row slices into every rank's buffer from the epilogue, barrier after them spans the GPUs" if sharded_plan else
                                "plan: one persistent kernel per step (per-warp TMA rings, in-kernel activation quantization, grid barrier between groups)" if len(segments) == 1 else
                                f"plan launches for the replicated groups ({sum(1 for sg in segments if sg[0] == 'plan')} launches) + fused-all-gather GEMVs for the sharded ones") if plan is not None
                               else "oneshot: one quantizer + one GEMV launch per group, CUDA graph"),
                    "cuda_graph": graphed, "split_groups": sum(1 for g2, _ in schedule if split_grp[g2]),
                    "replicated_groups": sum(1 for g2, _ in schedule if not split_grp[g2]), "split_min_mb": args.split_min_mb,
                    "parallelism": (f"{world} replicas of the single-GPU step (no launch group is worth an exchange at this size and rank count: DESIGN 6.1)" if not any(split_grp.values()) else f"row-shard x{world} (launch groups >= {args.split_min_mb} MB; smaller ones replicated) + " + ("fused all-gather (peer stores over NVLink + flag barrier per launch group); data plane is not NCCL" if pg is not None else "NCCL all-gather per launch group")) if world > 1 else "single GPU",
                    "decode_tokens_per_s": 1e3 / ms_per_step})
        line = {"metric": METRIC[wtype], "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "int8 dp4a block dots, fp32 accumulate", "data": "synthetic", "config": cfg,
                "roofline": roofline, "cpu_baseline": cpu_baseline, "e2e": e2e,
                "gpu_launches": int(launches_per_step * args.steps), "clocks": clocks, "parity_checked": parity_checked}
        real_stdout.write(json.dumps(line) + "\n")
        real_stdout.flush()
    log("line printed; leaving")
    # orderly exit (the driver's exit hook records which native libraries this process loaded): graphs first, then the
    # plan / peer group, then the process group
    graphs.clear()
    del run_full
    if plan is not None:
        plan.free()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    sys.stdout.flush()
    sys.stderr.flush()
    if timed_out:
        raise SystemExit("bench: a cross-GPU / grid barrier timed out; no result line printed")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="llama3-8b-q4_0-decode", choices=sorted(WORKLOADS))
    ap.add_argument("--decode", default="plan", choices=["plan", "oneshot"], help="plan = one persistent kernel per step (Q4_0, one GPU); oneshot = a launch per group")
    ap.add_argument("--layers", type=int, default=0, help="override the layer count (debug)")
    ap.add_argument("--no-e2e", dest="e2e", action="store_false")
    ap.add_argument("--e2e-pinned", action="store_true", help="e2e with pinned host B/C (default: pageable, as stock ggml allocates)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the prefill / mixmul / quantizer legs (roofline.prefill etc.)")
    ap.add_argument("--split-min-mb", type=float, default=None,
                    help="N>1: launch groups with fewer weight MB are computed on every rank instead of sharded (default: the break-even of the path in use)")
    ap.add_argument("--gather", default="p2p", choices=["p2p", "nccl"], help="multi-GPU output exchange")
    ap.add_argument("--no-graph", action="store_true", help="time eager launches instead of a CUDA graph replay (debug)")
    args = ap.parse_args()
    args.warmup = max(3, args.warmup) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()
