"""Generates tests/golden/hotpath_golden.npz by RUNNING THE REFERENCE's own compiled hot path
(oracle/_ref/libtb200_ref.so, built by oracle/build_ref.sh from /root/reference). The reference keeps no
golden vectors for llamafile_sgemm / llamafile_mixmul / the Q8 quantizers, so these
fixtures are how its behaviour travels to the GPU box, where /root/reference does not exist.

    python tests/golden/make_golden.py
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
from oracle import binding as B  # noqa: E402
import helpers as H  # noqa: E402


def main():
    B.build_ref()
    r = B.reference()
    assert r is not None, "reference library missing"
    g = {"variant": np.array(r.variant)}
    # --- quantizers: adversarial + random rows
    adv = H.adversarial_rows(256)
    rnd = np.random.default_rng(11).standard_normal((8, 1024)).astype(np.float32)
    g["q_adv_x"], g["q_rnd_x"] = adv, rnd
    g["q_adv_q8_0"], g["q_adv_q8_K"] = r.quantize_q8_0(adv), r.quantize_q8_K(adv)
    g["q_rnd_q8_0"], g["q_rnd_q8_K"] = r.quantize_q8_0(rnd), r.quantize_q8_K(rnd)
    g["q_rnd_f16"] = r.fp32_to_fp16(rnd)
    # --- llamafile_sgemm through ggml's mul_mat flow, every weight type, decode and small-batch shapes
    N, K = 48, 512
    for wt in H.WEIGHT_TYPES:
        W = {H.Q4_0: r.quantize_q4_0, H.Q8_0: r.quantize_q8_0_ref, H.Q4_K: r.quantize_q4_K, H.Q6_K: r.quantize_q6_K, H.F16: r.fp32_to_fp16}[wt](
            (np.random.default_rng(100 + wt).standard_normal((N, K)) * 0.02).astype(np.float32))
        g[f"w_{wt}"] = W
        for T in (1, 2, 3, 8, 9, 33):
            X = H.make_acts(T, K, seed=200 + T)
            g[f"x_{T}"] = X
            g[f"mm_{wt}_{T}"] = r.mul_mat(W, wt, X, nth=3)
    # TinyLLama-5M-like F16/Q8_0 layer shapes (config 1; the model file itself is not in the checkout)
    for (n, k) in ((64, 64), (256, 64), (64, 256)):
        Wf = (np.random.default_rng(n * 1000 + k).standard_normal((n, k)) * 0.05).astype(np.float32)
        X = H.make_acts(5, k, seed=n + k)
        g[f"tl_x_{n}_{k}"] = X
        for wt, q in ((H.F16, r.fp32_to_fp16), (H.Q8_0, r.quantize_q8_0_ref)):
            W = q(Wf)
            g[f"tl_w_{wt}_{n}_{k}"] = W
            g[f"tl_mm_{wt}_{n}_{k}"] = r.mul_mat(W, wt, X, nth=2)
            g[f"tl_mv_{wt}_{n}_{k}"] = r.mul_mat(W, wt, X[:1], nth=2)
    # --- llamafile_mixmul: 4 experts, top-2, with and without a shared task row
    rows, cols, experts, thinkers = 32, 256, 4, 2
    for wt in (H.Q4_0, H.Q8_0, H.F16):
        Wf = (np.random.default_rng(300 + wt).standard_normal((experts * rows, cols)) * 0.02).astype(np.float32)
        W = {H.Q4_0: r.quantize_q4_0, H.Q8_0: r.quantize_q8_0_ref, H.F16: r.fp32_to_fp16}[wt](Wf)
        g[f"mx_w_{wt}"] = W
        for tokens, tasks in ((1, 1), (5, 1), (5, 2), (19, 2)):
            X = np.random.default_rng(400 + tokens + tasks).standard_normal((tokens, tasks, cols)).astype(np.float32)
            plan = H.make_plan(tokens, experts, thinkers, seed=500 + tokens)
            out = np.full((tokens, thinkers, rows), np.nan, np.float32)
            ok = r.mixmul(B.Tensor.of(W, wt, [cols, rows, experts]), B.Tensor.of(X, H.F32, [cols, tasks, tokens]),
                          B.Tensor.of(plan, H.I32, [thinkers, tokens]), B.Tensor.of(out, H.F32, [rows, thinkers, tokens]), nth=3)
            assert ok
            g[f"mx_x_{tokens}_{tasks}"], g[f"mx_plan_{tokens}"] = X, plan
            g[f"mx_out_{wt}_{tokens}_{tasks}"] = out
    # --- llamafile_mixmul_iqk (Q4_K experts): one expert's rows with a row mapping
    Nx, ne00, ne11, ne12 = 40, 512, 2, 6
    W = r.quantize_q4_K((np.random.default_rng(600).standard_normal((Nx, ne00)) * 0.02).astype(np.float32))
    X = np.random.default_rng(601).standard_normal((ne12 * ne11, ne00)).astype(np.float32)
    Bq = r.quantize_q8_K(X)
    mapping = np.array([[0, 0], [1, 2], [0, 3], [1, 5], [1, 0]], np.int32)
    out = np.full((ne12, 2, Nx), np.nan, np.float32)
    assert r.mixmul_iqk(Nx, len(mapping), ne00, ne11, H.Q4_K, W, Bq, out, Nx * 4, 2 * Nx * 4, mapping, nth=2)
    g["iqk_w"], g["iqk_x"], g["iqk_map"], g["iqk_out"] = W, X, mapping, out
    path = ROOT / "tests" / "golden" / "hotpath_golden.npz"
    np.savez_compressed(path, **g)
    print(path, path.stat().st_size, "bytes;", len(g), "arrays; reference variant", r.variant)


if __name__ == "__main__":
    main()
