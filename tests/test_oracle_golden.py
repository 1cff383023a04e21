"""CPU: the oracle (oracle/tb_oracle.c) against the committed fixtures that were produced by running the
reference's own compiled hot path (tests/golden/make_golden.py). Bit-exact for the quantizers, scaled
error for fp32 outputs."""
from pathlib import Path

import numpy as np
import pytest

import helpers as H
from oracle import binding as B

ROOT = Path(__file__).resolve().parent.parent


def test_quantizers_bit_exact(oracle, golden):
    for tag in ("adv", "rnd"):
        x = golden[f"q_{tag}_x"]
        assert np.array_equal(oracle.quantize_q8_0(x), golden[f"q_{tag}_q8_0"]), tag
        assert np.array_equal(oracle.quantize_q8_K(x), golden[f"q_{tag}_q8_K"]), tag
    assert np.array_equal(oracle.fp32_to_fp16(golden["q_rnd_x"]), golden["q_rnd_f16"])


@pytest.mark.parametrize("wt", H.WEIGHT_TYPES)
@pytest.mark.parametrize("T", [1, 2, 3, 8, 9, 33])
def test_mul_mat(oracle, golden, wt, T):
    out = oracle.mul_mat(golden[f"w_{wt}"], wt, golden[f"x_{T}"])
    H.assert_close(out, golden[f"mm_{wt}_{T}"], f"{H.TYPE_NAME[wt]} T={T}")
    assert H.scaled_err(out, golden[f"mm_{wt}_{T}"]) < 5e-6   # the oracle itself sits within accumulation noise


@pytest.mark.parametrize("shape", [(64, 64), (256, 64), (64, 256)])
@pytest.mark.parametrize("wt", [H.F16, H.Q8_0])
def test_tinyllama_shapes(oracle, golden, wt, shape):
    n, k = shape
    X = golden[f"tl_x_{n}_{k}"]
    H.assert_close(oracle.mul_mat(golden[f"tl_w_{wt}_{n}_{k}"], wt, X), golden[f"tl_mm_{wt}_{n}_{k}"])
    H.assert_close(oracle.mul_mat(golden[f"tl_w_{wt}_{n}_{k}"], wt, X[:1]), golden[f"tl_mv_{wt}_{n}_{k}"])


@pytest.mark.parametrize("wt", [H.Q4_0, H.Q8_0, H.F16])
@pytest.mark.parametrize("tokens,tasks", [(1, 1), (5, 1), (5, 2), (19, 2)])
def test_mixmul(oracle, golden, wt, tokens, tasks):
    rows, cols, experts, thinkers = 32, 256, 4, 2
    W, X, plan = golden[f"mx_w_{wt}"], golden[f"mx_x_{tokens}_{tasks}"], golden[f"mx_plan_{tokens}"]
    out = np.full((tokens, thinkers, rows), np.nan, np.float32)
    assert oracle.mixmul(B.Tensor.of(W, wt, [cols, rows, experts]), B.Tensor.of(X, H.F32, [cols, tasks, tokens]),
                         B.Tensor.of(plan, H.I32, [thinkers, tokens]), B.Tensor.of(out, H.F32, [rows, thinkers, tokens]))
    H.assert_close(out, golden[f"mx_out_{wt}_{tokens}_{tasks}"])


def test_mixmul_iqk(oracle, golden):
    W, X, mapping = golden["iqk_w"], golden["iqk_x"], golden["iqk_map"]
    Nx, ne00, ne11, ne12 = 40, 512, 2, 6
    out = np.full((ne12, 2, Nx), np.nan, np.float32)
    assert oracle.mixmul_iqk(Nx, len(mapping), ne00, ne11, H.Q4_K, W, oracle.quantize_q8_K(X), out, Nx * 4, 2 * Nx * 4, mapping)
    ref = golden["iqk_out"]
    assert np.array_equal(np.isnan(out), np.isnan(ref))   # rows no mapping names stay untouched
    H.assert_close(out[~np.isnan(ref)], ref[~np.isnan(ref)])


def test_block_dots_match_vec_dot(oracle):
    """int32 block sums recombine to the oracle's own vec_dot (exactness of the integer part)."""
    K = 1024
    x = H.make_acts(1, K, 5)
    for wt in (H.Q4_0, H.Q8_0, H.Q4_K):
        W = H.make_weights(wt, 1, K, seed=9, realistic_q4k=False)
        Y = oracle.from_float(x, B.VEC_DOT_TYPE[wt])
        sums = oracle.blockdot(wt, W, Y, K // B.BLCK[wt])
        assert sums.dtype == np.int32 and np.abs(sums).max() > 0
        if wt != H.Q4_K:
            wd = np.frombuffer(W.tobytes(), np.uint8).reshape(-1, B.TYPE_SIZE[wt])[:, :2].copy().view(np.float16).astype(np.float32).ravel()
            yd = np.frombuffer(Y.tobytes(), np.uint8).reshape(-1, 34)[:, :2].copy().view(np.float16).astype(np.float32).ravel()
            acc = np.float32(0)
            for s, a, b in zip(sums, wd, yd):
                acc = np.float32(acc + np.float32(np.float32(np.float32(s) * a) * b))
            assert acc == np.float32(oracle.vec_dot(wt, K, W, Y))


def test_fused_layer_golden(oracle):
    """The oracle reproduces the reference's one-layer decode chain (matmuls + rms_norm / silu / mul / add neighbours),
    weights regenerated from the generator's seeds with the oracle's quantizer (CRC proves they are the reference's bytes)."""
    import sys
    import zlib
    sys.path.insert(0, str(ROOT / "tests" / "golden"))
    import make_golden_fused as M
    g = np.load(ROOT / "tests" / "golden" / "fused_golden.npz")
    W = M.layer_weights(oracle.quantize_q4_0)
    assert [zlib.crc32(W[k].tobytes()) for k in M.SHAPES] == list(g["crc"])
    I = {k: g[f"in_{k}"] for k in ("x", "attn", "norm1", "norm2")}
    got = M.run_layer(oracle, W, I)
    assert np.array_equal(got["h1"], g["out_h1"])            # rms_norm * weight: bit-exact
    for k, v in got.items():
        H.assert_close(v[None], g[f"out_{k}"][None], k)


@pytest.mark.parametrize("wt", H.G8_TYPES)
def test_g8_types_golden(oracle, wt):
    """SURVEY.md §8 f3 types: the oracle's dequantizers are bit-equal to the reference's, its products agree with llamafile_sgemm."""
    g = np.load(ROOT / "tests" / "golden" / "g8_golden.npz")
    W = g[f"w_{wt}"]
    assert np.array_equal(oracle.dequantize(wt, W, 40 * 512), g[f"deq_{wt}"])
    for T in (1, 3, 8, 33):
        H.assert_close(oracle.mul_mat(W, wt, g[f"x_{wt}_{T}"]), g[f"mm_{wt}_{T}"], f"{H.TYPE_NAME[wt]} T={T}")
    assert np.array_equal(oracle.quantize_q8_1(g["q81_x"]), g["q81_y"])
