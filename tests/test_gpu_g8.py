"""GPU (B200): the remaining weight types iqk serves on the same entry points (llamafile/iqk_mul_mat.inc:1425-1447)
— Q4_1 (Q8_1 activations), Q5_0 (Q8_0), Q5_K and IQ4_XS (Q8_K) — held in the generic resident form G8 (int8 quants + fp16 {a, b}
per 32-block, tb200_common.cuh): decode kernels, tcgen05 prefill, the host entry points and the ggml backend."""
from pathlib import Path

import numpy as np
import pytest

import helpers as H
from oracle import binding as B

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")
ROOT = Path(__file__).resolve().parent.parent


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def gpu_mul_mat(tb, W, wt, X, rows, k):
    w = tb.Weights(W, wt, rows, k)
    out = torch.full((X.shape[0], rows), float("nan"), device="cuda", dtype=torch.float32)
    tb.mul_mat(w, dev(X), X.shape[0], out)
    tb.synchronize()
    res = out.cpu().numpy()
    w.free()
    return res


@pytest.mark.parametrize("wt", H.G8_TYPES)
@pytest.mark.parametrize("T", [1, 2, 3, 5, 8])
@pytest.mark.parametrize("N,K", [(64, 256), (131, 1024), (5000, 512)])
def test_g8_decode_parity(tb, oracle, wt, T, N, K):
    W = H.make_weights(wt, N, K, seed=N + T, realistic_q4k=False)
    X = H.make_acts(T, K, seed=T)
    H.assert_close(gpu_mul_mat(tb, W, wt, X, N, K), oracle.mul_mat(W, wt, X), f"{H.TYPE_NAME[wt]} N={N} K={K} T={T}")


@pytest.mark.parametrize("wt", H.G8_TYPES)
@pytest.mark.parametrize("N,K,T", [(128, 256, 32), (130, 512, 33), (384, 4096, 512), (1000, 2048, 300), (256, 512, 9)])
def test_g8_prefill_parity(tb, oracle, wt, N, K, T):
    W = H.make_weights(wt, N, K, seed=N + T, realistic_q4k=False)
    X = H.make_acts(T, K, seed=T)
    H.assert_close(gpu_mul_mat(tb, W, wt, X, N, K), oracle.mul_mat(W, wt, X), f"prefill {H.TYPE_NAME[wt]} N={N} K={K} T={T}")


@pytest.mark.parametrize("wt", H.G8_TYPES)
@pytest.mark.parametrize("T", [1, 3, 8, 33])
def test_g8_golden(tb, wt, T):
    """Against the fixtures the reference produced (weights from its quantize_row_*_ref; tests/golden/make_golden_g8.py)."""
    g = np.load(ROOT / "tests" / "golden" / "g8_golden.npz")
    out = gpu_mul_mat(tb, g[f"w_{wt}"], wt, g[f"x_{wt}_{T}"], 40, 512)
    H.assert_close(out, g[f"mm_{wt}_{T}"], f"golden {H.TYPE_NAME[wt]} T={T}")


@pytest.mark.parametrize("wt", H.G8_TYPES)
def test_g8_exact_integer_arithmetic(tb, oracle, wt):
    """Power-of-two scales and activations that quantize to themselves: the repack (5th bits, 6-bit sub-block scales, IQ4 codebook,
    mins) and the int8 block dots must reproduce the oracle bit for bit."""
    N, K, T = 96, 1024, 2
    rng = np.random.default_rng(wt)
    ts, blck = B.TYPE_SIZE[wt], B.BLCK[wt]
    b = rng.integers(0, 256, (N * K // blck, ts), dtype=np.uint8)
    b[:, 0:2] = np.full(len(b), 2.0 ** -7, np.float16).view(np.uint8).reshape(-1, 2)
    if wt in (H.Q4_1, H.Q5_K):
        b[:, 2:4] = np.full(len(b), 2.0 ** -6, np.float16).view(np.uint8).reshape(-1, 2)
    if wt == H.IQ4_XS:   # keep (ls - 32) * d a power of two too: ls = 33 for every sub-block (scales_l nibbles = 1, scales_h pairs = 2)
        b[:, 2:4] = np.array([0xAA, 0xAA], np.uint8)
        b[:, 4:8] = 0x11
    if wt == H.Q5_K:     # sub-block scales 1..., mins: any 6-bit values are exact with power-of-two d / dmin
        pass
    W = b.reshape(-1)
    X = rng.integers(-6, 7, (T, K)).astype(np.float32)
    X[:, ::blck] = 127.0
    out = gpu_mul_mat(tb, W, wt, X, N, K)
    ref = oracle.mul_mat(W, wt, X)
    assert np.array_equal(out, ref), float(np.abs(out - ref).max())


@pytest.mark.parametrize("wt", H.G8_TYPES)
@pytest.mark.parametrize("n", [1, 4, 40])
def test_g8_llamafile_sgemm_host(tb, oracle, wt, n):
    m, K = 200, 1024
    W = H.make_weights(wt, m, K, seed=n, realistic_q4k=False)
    X = H.make_acts(n, K, seed=10 + n)
    kb = K // B.BLCK[wt]
    vdt = B.VEC_DOT_TYPE[wt]
    Bq = oracle.from_float(X, vdt)
    Cm = np.full((n, m), np.nan, np.float32)
    assert tb.llamafile_sgemm(m, n, kb, W, kb, Bq, kb, Cm, m, 0, 1, wt, vdt, H.F32)
    ok, Cr = oracle.sgemm(m, n, kb, W, kb, Bq, kb, m, wt, vdt)
    assert ok
    H.assert_close(Cm, Cr, f"{H.TYPE_NAME[wt]} n={n}")
    # the wrong activation type is "not handled", C untouched
    wrong = H.Q8_0 if vdt != H.Q8_0 else H.Q8_K
    C0 = np.full((n, m), np.nan, np.float32)
    assert not tb.llamafile_sgemm(m, n, kb, W, kb, Bq, kb, C0, m, 0, 1, wt, wrong, H.F32)
    assert np.isnan(C0).all()


@pytest.mark.parametrize("typeA", H.G8_TYPES)
def test_g8_llamafile_mixmul_iqk_host(tb, oracle, typeA):
    Nx, ne00, ne11, ne12 = 24, 512, 2, 9
    W = H.make_weights(typeA, Nx, ne00, seed=3, realistic_q4k=False)
    X = np.random.default_rng(4).standard_normal((ne12 * ne11, ne00)).astype(np.float32)
    mapping = np.array([[i % 2, (3 * i) % ne12] for i in range(11)], np.int32)
    Bq = oracle.from_float(X, B.VEC_DOT_TYPE[typeA])
    out = np.full((ne12, 2, Nx), np.nan, np.float32)
    assert tb.llamafile_mixmul_iqk(Nx, len(mapping), ne00, ne11, typeA, W, Bq, out, Nx * 4, 2 * Nx * 4, mapping, 0, 1)
    ref = np.full((ne12, 2, Nx), np.nan, np.float32)
    assert oracle.mixmul_iqk(Nx, len(mapping), ne00, ne11, typeA, W, Bq, ref, Nx * 4, 2 * Nx * 4, mapping)
    assert np.array_equal(np.isnan(out), np.isnan(ref))
    m = ~np.isnan(ref)
    H.assert_close(out[m], ref[m])


@pytest.mark.parametrize("wt", H.G8_TYPES)
def test_g8_backend_mul_mat(tb, wt):
    """Through the ggml backend object, against ggml's own CPU backend (its vec_dot of the type)."""
    from ggml_host import Graph, ggml
    G = ggml()
    if G is None:
        pytest.skip("oracle/_ref not built")
    assert tb.backend_link(G.ref_backend_api())
    backend, cpu = tb.backend_init(0), G.ggml_backend_cpu_init()
    N, K, T = 130, 1024, 5
    W = H.make_weights(wt, N, K, seed=2, realistic_q4k=False)
    X = H.make_acts(T, K, seed=3)
    res = []
    for bk in (backend, cpu):
        g = Graph(bk)
        a, b = g.tensor(wt, K, N), g.tensor(H.F32, K, T)
        out = g.mul_mat(a, b)
        g.build([out])
        g.set(a, W); g.set(b, X)
        assert g.compute() == 0
        res.append(g.get(out, (T, N)))
        g.free()
    H.assert_close(res[0], res[1], f"backend {H.TYPE_NAME[wt]}")
    G.ggml_backend_free(backend); G.ggml_backend_free(cpu)
