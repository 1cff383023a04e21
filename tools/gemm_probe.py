#!/usr/bin/env python
"""Sustained prefill-GEMM probe: runs one shape back to back for --seconds while sampling SM clock and board power
(NVML), next to torch.matmul fp16 (cuBLAS) on the same shape. Answers "is the kernel clock/power limited?".
    python tools/gemm_probe.py --type q4_0 --N 14336 --K 4096 --T 4096
"""
import argparse, json, sys, threading, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tools"))
import torch
import pynvml
from bench_prefill import TYPES, synth
import llamafile_b200 as tb


class Sampler(threading.Thread):
    def __init__(self):
        super().__init__(daemon=True)
        pynvml.nvmlInit()
        self.h = pynvml.nvmlDeviceGetHandleByIndex(torch.cuda.current_device())
        self.rows, self.stop = [], False
    def run(self):
        while not self.stop:
            self.rows.append((pynvml.nvmlDeviceGetClockInfo(self.h, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetPowerUsage(self.h) / 1e3))
            time.sleep(0.02)
    def summary(self):
        r = self.rows[len(self.rows) // 3:]   # skip ramp
        if not r: return {}
        c = sorted(x[0] for x in r); pw = sorted(x[1] for x in r)
        return {"sm_mhz_median": c[len(c) // 2], "sm_mhz_min": c[0], "power_w_median": pw[len(pw) // 2], "power_w_max": pw[-1], "samples": len(r)}


def sustained(fn, stream, seconds, flops):
    with torch.cuda.stream(stream):
        for _ in range(3): fn()
        stream.synchronize()
        s = Sampler(); s.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n, t0 = 0, time.perf_counter()
        e0.record(stream)
        while time.perf_counter() - t0 < seconds:
            for _ in range(20): fn()
            n += 20
            stream.synchronize()
        e1.record(stream); stream.synchronize()
        s.stop = True; s.join()
    ms = e0.elapsed_time(e1) / n
    return {"ms": ms, "TFLOPS": flops / (ms * 1e-3) / 1e12, **s.summary()}


ap = argparse.ArgumentParser()
ap.add_argument("--type", default="q4_0"); ap.add_argument("--N", type=int, default=14336); ap.add_argument("--K", type=int, default=4096)
ap.add_argument("--T", type=int, default=4096); ap.add_argument("--seconds", type=float, default=1.5); ap.add_argument("--cublas", type=int, default=1)
a = ap.parse_args()
assert tb.link()
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(); tb.set_stream(stream.cuda_stream)
fl = 2.0 * a.T * a.N * a.K
for name in a.type.split(","):
    w = tb.Weights(synth(torch, dev, name, a.N, a.K), TYPES[name][0], a.N, a.K)
    x = torch.randn(a.T, a.K, device=dev); out = torch.empty(a.T, a.N, device=dev)
    acts = tb.Acts(TYPES[name][0], a.K, a.T)
    with torch.cuda.stream(stream):
        acts.quantize(x, a.T)
    r = sustained(lambda: tb.mul_mat_acts(w, acts, out), stream, a.seconds, fl)
    print(json.dumps({"impl": "tb200", "type": name, "N": a.N, "K": a.K, "T": a.T, "includes": "fp16 prep + GEMM (activations pre-quantized)", **r}), flush=True)
    w.free()
if a.cublas:
    A = torch.randn(a.N, a.K, device=dev, dtype=torch.float16); Bm = torch.randn(a.T, a.K, device=dev, dtype=torch.float16)
    r = sustained(lambda: torch.matmul(Bm, A.t()), stream, a.seconds, fl)
    print(json.dumps({"impl": "cublas fp16", "N": a.N, "K": a.K, "T": a.T, **r}), flush=True)
