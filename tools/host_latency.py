#!/usr/bin/env python
"""Per-call latency of llamafile_sgemm() on host buffers (decode, n=1), Q4_0.
    python tools/host_latency.py [--n 4096,14336,128256] [--pinned 0|1]
"""
import argparse, json, sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np
import torch
import llamafile_b200 as tb
from llamafile_b200.lib import lib, Q4_0, Q8_0, F32

ap = argparse.ArgumentParser()
ap.add_argument("--n", default="1024,4096,14336,128256")
ap.add_argument("--pinned", type=int, default=1)
ap.add_argument("--iters", type=int, default=2000)
a = ap.parse_args()
assert tb.link()
s = torch.cuda.Stream(); tb.set_stream(s.cuda_stream)
K = 4096; kb = K // 32
fn = lib().llamafile_sgemm
rng = np.random.default_rng(0)
for N in [int(v) for v in a.n.split(",")]:
    W = rng.integers(0, 256, N * kb * 18, dtype=np.uint8)
    Wb = W.reshape(-1, 18); Wb[:, :2] = np.full(Wb.shape[0], 1e-3, np.float16).view(np.uint8).reshape(-1, 2)
    x = rng.standard_normal(K).astype(np.float32)
    q = tb.quantize_row_q8_0(x)
    hB = torch.from_numpy(q); hC = torch.empty(N)
    if a.pinned:
        hB, hC = hB.pin_memory(), hC.pin_memory()
    args = (N, 1, kb, W.ctypes.data, kb, hB.data_ptr(), kb, hC.data_ptr(), N, 0, 1, Q4_0, Q8_0, F32)
    for _ in range(20):
        assert fn(*args)
    it = a.iters if N < 100000 else a.iters // 4
    t0 = time.perf_counter()
    for _ in range(it):
        fn(*args)
    dt = (time.perf_counter() - t0) / it
    print(json.dumps({"N": N, "K": K, "pinned": bool(a.pinned), "us_per_call": dt * 1e6}), flush=True)
    tb.weights_cache_clear()
