"""Generates tests/golden/fused_golden.npz by RUNNING THE REFERENCE (oracle/_ref/libtb200_ref.so: ggml's own CPU ops for
rms_norm / silu / mul / add, and llamafile_sgemm for the matmuls) on one transformer layer's worth of decode matmuls with
their graph neighbours (llm_build_norm, llm_build_ffn, residual adds: llama.cpp/src/llama.cpp:8099-8225). Weights are NOT
stored: the tests regenerate them from the seeds below with the oracle's quantize_row_q4_0_ref, which is pinned bit-exact
to the reference's (tests/test_oracle_vs_reference.py); a CRC of the reference-quantized bytes is stored to prove it.

    python tests/golden/make_golden_fused.py
"""
import sys
import zlib
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
from oracle import binding as B  # noqa: E402

HID, KV, FFN, EPS = 512, 128, 1280, 1e-5
SHAPES = {"wq": (HID, HID), "wk": (KV, HID), "wv": (KV, HID), "wo": (HID, HID), "gate": (FFN, HID), "up": (FFN, HID), "down": (HID, FFN)}


def layer_weights(quantize):
    out = {}
    for i, (name, (n, k)) in enumerate(SHAPES.items()):
        w = (np.random.default_rng(900 + i).standard_normal((n, k)) * (2.0 / np.sqrt(k))).astype(np.float32)
        out[name] = quantize(w)
    return out


def layer_inputs():
    rng = np.random.default_rng(77)
    return {"x": rng.standard_normal(HID).astype(np.float32) * 1.5, "attn": rng.standard_normal(HID).astype(np.float32),
            "norm1": (1 + 0.2 * rng.standard_normal(HID)).astype(np.float32), "norm2": (1 + 0.2 * rng.standard_normal(HID)).astype(np.float32)}


def run_layer(impl, W, I):
    """The decode matmuls of one layer with their neighbours, as the reference graph orders them."""
    mm = lambda w, v: impl.mul_mat(W[w], B.Q4_0, v[None])[0]   # noqa: E731
    g = {}
    g["h1"] = impl.rms_norm_mul(I["x"], I["norm1"], EPS)
    for w in ("wq", "wk", "wv"):
        g[w] = mm(w, g["h1"])
    g["x2"] = impl.add(mm("wo", I["attn"]), I["x"]) if hasattr(impl, "add") else (mm("wo", I["attn"]) + I["x"]).astype(np.float32)
    g["h2"] = impl.rms_norm_mul(g["x2"], I["norm2"], EPS)
    g["gate"], g["up"] = mm("gate", g["h2"]), mm("up", g["h2"])
    g["act"] = impl.silu_mul(g["gate"], g["up"])
    d = mm("down", g["act"])
    g["x3"] = impl.add(d, g["x2"]) if hasattr(impl, "add") else (d + g["x2"]).astype(np.float32)
    return g


def main():
    B.build_ref()
    r = B.reference()
    assert r is not None, "reference library missing"
    W = layer_weights(r.quantize_q4_0)
    I = layer_inputs()
    g = {f"in_{k}": v for k, v in I.items()}
    g.update({f"out_{k}": v for k, v in run_layer(r, W, I).items()})
    g["crc"] = np.array([zlib.crc32(W[k].tobytes()) for k in SHAPES], np.uint32)
    g["variant"] = np.array(r.variant)
    np.savez_compressed(ROOT / "tests" / "golden" / "fused_golden.npz", **g)
    print("wrote fused_golden.npz:", {k: v.shape for k, v in g.items()})


if __name__ == "__main__":
    main()
