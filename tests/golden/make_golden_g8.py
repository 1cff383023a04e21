"""Generates tests/golden/g8_golden.npz by RUNNING THE REFERENCE (oracle/_ref) for the SURVEY.md §8 f3 weight types — Q4_1, Q5_0,
Q5_K, IQ4_XS — that iqk serves on the same entry points (llamafile/iqk_mul_mat.inc:1425-1447): weights quantized by the
reference's quantize_row_*_ref, activations through its from_float of the vec_dot type (Q8_1 / Q8_0 / Q8_K), products through
llamafile_sgemm.

IQ4_XS note: the reference's IQ4_XS kernels add 128 to the codebook values and accumulate products of those unsigned bytes in
int16 lanes with saturation (iqk_mul_mat.inc:432-470 zen4, :770-800 avx2); with activation quants of typical Gaussian size the
lanes saturate and the reference deviates from ggml's own ggml_vec_dot_iq4_xs_q8_K by 5-20 %. The fixtures therefore use
activations with one large element per 256-block (what real post-norm activations look like), where the kernels are exact.

    python tests/golden/make_golden_g8.py
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
from oracle import binding as B  # noqa: E402
import helpers as H  # noqa: E402


def acts(wt, T, K, seed):
    X = H.make_acts(T, K, seed=seed)
    if wt == H.IQ4_XS:
        X[:, ::256] = 60.0   # keeps the Q8_K quants of the other elements small (see the IQ4_XS note)
    return X


def main():
    B.build_ref()
    r = B.reference()
    assert r is not None
    g = {"variant": np.array(r.variant)}
    N, K = 40, 512
    for wt in H.G8_TYPES:
        W = r.quantize_type(wt, (np.random.default_rng(700 + wt).standard_normal((N, K)) * 0.02).astype(np.float32))
        g[f"w_{wt}"] = W
        g[f"deq_{wt}"] = r.dequantize(wt, W, N * K)
        for T in (1, 3, 8, 33):
            X = acts(wt, T, K, 800 + T)
            g[f"x_{wt}_{T}"] = X
            g[f"mm_{wt}_{T}"] = r.mul_mat(W, wt, X, nth=2)
    x = H.make_acts(4, 256, seed=5)
    g["q81_x"], g["q81_y"] = x, r.quantize_q8_1(x)
    np.savez_compressed(ROOT / "tests" / "golden" / "g8_golden.npz", **g)
    print("wrote g8_golden.npz", len(g), "arrays")


if __name__ == "__main__":
    main()
