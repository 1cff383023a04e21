"""CPU, this container only: the oracle against the reference's own compiled code (oracle/_ref), on fresh
seeded inputs and at larger shapes than the fixtures. Skipped where /root/reference is absent."""
import ctypes as C

import numpy as np
import pytest

import helpers as H
from oracle import binding as B


def test_reference_variant(reference):
    assert reference.variant in ("amd_zen4", "amd_avx512f", "amd_avx2")


@pytest.mark.parametrize("k", [32, 256, 4096])
def test_quantizers_bit_exact(oracle, reference, k):
    rng = np.random.default_rng(k)
    for scale in (1.0, 1e-3, 300.0):
        x = (rng.standard_normal((33, k)) * scale).astype(np.float32)
        assert np.array_equal(oracle.quantize_q8_0(x), reference.quantize_q8_0(x))
        assert np.array_equal(oracle.quantize_q8_0_ref(x), reference.quantize_q8_0_ref(x))
        assert np.array_equal(oracle.quantize_q4_0(x), reference.quantize_q4_0(x))
        assert np.array_equal(oracle.fp32_to_fp16(x), reference.fp32_to_fp16(x))
        if k % 256 == 0:
            assert np.array_equal(oracle.quantize_q8_K(x), reference.quantize_q8_K(x))
    adv = H.adversarial_rows(max(k, 256))
    assert np.array_equal(oracle.quantize_q8_0(adv), reference.quantize_q8_0(adv))
    assert np.array_equal(oracle.quantize_q8_K(adv), reference.quantize_q8_K(adv))


def test_runtime_q8_0_differs_from_ref_body(reference):
    """the runtime AVX body and quantize_row_q8_0_ref are not bit-identical; the target is the runtime one."""
    x = np.random.default_rng(0).standard_normal((512, 4096)).astype(np.float32)
    a, b = reference.quantize_q8_0(x), reference.quantize_q8_0_ref(x)
    assert 0 < (a != b).sum() < 1000


@pytest.mark.parametrize("wt", H.WEIGHT_TYPES)
@pytest.mark.parametrize("N,K,T", [(64, 256, 1), (130, 1024, 4), (96, 4096, 8), (70, 512, 17)])
def test_sgemm_flow(oracle, reference, wt, N, K, T):
    W = H.make_weights(wt, N, K, seed=N + K)
    X = H.make_acts(T, K, seed=T)
    a, b = oracle.mul_mat(W, wt, X), reference.mul_mat(W, wt, X, nth=4)
    H.assert_close(a, b, f"{H.TYPE_NAME[wt]} {N}x{K}x{T}")
    assert H.scaled_err(a, b) < 5e-6


def test_sgemm_service_matrix(oracle, reference):
    """Which (Atype, Btype, n) are serviced (true) vs handed back to ggml (false) must agree."""
    K, N = 256, 8
    X = H.make_acts(3, K)
    for wt in H.WEIGHT_TYPES:
        W = H.make_weights(wt, N, K)
        kb = K // B.BLCK[wt]
        # pairings a ggml caller can produce: B is F32 (first attempt), F16, or A's vec_dot type. (Q4_K with Q8_0 or
        # Q4_0 with Q8_K violate the reference's own asserts / reinterpret bytes and are outside the contract.)
        for bt in sorted({H.F32, H.F16, B.VEC_DOT_TYPE[wt]}):
            if bt == H.F16 and wt != H.F16:
                continue
            Bq = X if bt == H.F32 else oracle.from_float(X, bt)
            ldb = K // B.BLCK[bt] if bt != H.F32 else K
            if bt in (H.F32, H.F16) and wt != H.F16:
                ldb = kb   # element-typed B against block-counted k: the call only has to be rejected
            for n in (1, 2, 3):
                ok_o, _ = oracle.sgemm(N, n, kb, W, kb, Bq, ldb, N, wt, bt)
                ok_r, _ = reference.sgemm(N, n, kb, W, kb, Bq, ldb, N, wt, bt)
                assert ok_o == ok_r, (wt, bt, n)


@pytest.mark.parametrize("wt", [H.Q4_0, H.Q8_0, H.F16])
@pytest.mark.parametrize("tokens,tasks,skew", [(1, 1, False), (7, 1, False), (7, 2, False), (40, 2, True)])
def test_mixmul(oracle, reference, wt, tokens, tasks, skew):
    rows, cols, experts, thinkers = 48, 512, 8, 2
    W = H.make_weights(wt, experts * rows, cols, seed=tokens)
    X = np.random.default_rng(tokens).standard_normal((tokens, tasks, cols)).astype(np.float32)
    plan = H.make_plan(tokens, experts, thinkers, skew=skew)
    outs = []
    for impl in (oracle, reference):
        out = np.full((tokens, thinkers, rows), np.nan, np.float32)
        tw, tx = B.Tensor.of(W, wt, [cols, rows, experts]), B.Tensor.of(X, H.F32, [cols, tasks, tokens])
        tp, to = B.Tensor.of(plan, H.I32, [thinkers, tokens]), B.Tensor.of(out, H.F32, [rows, thinkers, tokens])
        assert impl.mixmul(tw, tx, tp, to, nth=3)
        outs.append(out)
    assert oracle.mixmul_needs(tw, tx, tp) == reference.mixmul_needs(tw, tx, tp)
    H.assert_close(outs[0], outs[1])


def test_mixmul_out_of_range_expert_untouched(oracle, reference):
    rows, cols, experts, thinkers, tokens = 16, 256, 4, 2, 3
    W = H.make_weights(H.Q8_0, experts * rows, cols)
    X = np.random.default_rng(1).standard_normal((tokens, 1, cols)).astype(np.float32)
    plan = np.array([[0, 1], [9, 2], [3, -1]], np.int32)
    outs = []
    for impl in (oracle, reference):
        out = np.full((tokens, thinkers, rows), 7.5, np.float32)
        assert impl.mixmul(B.Tensor.of(W, H.Q8_0, [cols, rows, experts]), B.Tensor.of(X, H.F32, [cols, 1, tokens]),
                           B.Tensor.of(plan, H.I32, [thinkers, tokens]), B.Tensor.of(out, H.F32, [rows, thinkers, tokens]))
        outs.append(out)
    assert (outs[0][1, 0] == 7.5).all() and (outs[0][2, 1] == 7.5).all()
    assert np.array_equal(outs[0] == 7.5, outs[1] == 7.5)
    H.assert_close(outs[0], outs[1])


def test_mixmul_q4_K_not_serviced(oracle, reference):
    """Q4_K experts go through llamafile_mixmul_iqk, not llamafile_mixmul (ggml.c.patch:1960-2018)."""
    rows, cols, experts = 8, 256, 2
    W = H.synth_q4_K(experts * rows, cols, 1)
    X = np.zeros((1, 1, cols), np.float32); plan = np.zeros((1, 1), np.int32); out = np.zeros((1, 1, rows), np.float32)
    for impl in (oracle, reference):
        assert not impl.mixmul(B.Tensor.of(W, H.Q4_K, [cols, rows, experts]), B.Tensor.of(X, H.F32, [cols, 1, 1]),
                               B.Tensor.of(plan, H.I32, [1, 1]), B.Tensor.of(out, H.F32, [rows, 1, 1]))


@pytest.mark.parametrize("typeA", [H.Q4_K, H.Q4_0, H.Q6_K])
def test_mixmul_iqk(oracle, reference, typeA):
    Nx, ne00, ne11, ne12 = 24, 512, 2, 9
    W = H.make_weights(typeA, Nx, ne00, seed=3)
    X = np.random.default_rng(4).standard_normal((ne12 * ne11, ne00)).astype(np.float32)
    Bq = oracle.from_float(X, B.VEC_DOT_TYPE[typeA])
    mapping = np.array([[i % 2, (3 * i) % ne12] for i in range(11)], np.int32)
    outs = []
    for impl in (oracle, reference):
        out = np.full((ne12, 2, Nx), np.nan, np.float32)
        assert impl.mixmul_iqk(Nx, len(mapping), ne00, ne11, typeA, W, Bq, out, Nx * 4, 2 * Nx * 4, mapping, nth=2)
        outs.append(out)
    assert np.array_equal(np.isnan(outs[0]), np.isnan(outs[1]))
    m = ~np.isnan(outs[1])
    H.assert_close(outs[0][m], outs[1][m])


# ------------------------------------------------------------------ the matmul's graph neighbours
@pytest.mark.parametrize("n", [256, 4096, 14336])
def test_rms_norm_mul_bit_exact_and_silu_mul_close(oracle, reference, n):
    rng = np.random.default_rng(n)
    x = (rng.standard_normal(n) * 3).astype(np.float32)
    w = (1 + 0.1 * rng.standard_normal(n)).astype(np.float32)
    assert np.array_equal(oracle.rms_norm_mul(x, w, 1e-5), reference.rms_norm_mul(x, w, 1e-5))
    g, u = (rng.standard_normal(n) * 2).astype(np.float32), rng.standard_normal(n).astype(np.float32)
    a, b = oracle.silu_mul(g, u), reference.silu_mul(g, u)
    # ggml's SIMD silu evaluates expf by a polynomial (ggml.c:2432-2594): a few ulp from the scalar definition
    assert float(np.abs(a - b).max() / np.abs(b).max()) < 1e-6


@pytest.mark.parametrize("tokens,used,rows", [(1, 2, 4096), (7, 2, 512), (5, 1, 64), (3, 4, 1000)])
def test_moe_combine_bit_exact(oracle, reference, tokens, used, rows):
    rng = np.random.default_rng(tokens + rows)
    e = rng.standard_normal((tokens, used, rows)).astype(np.float32)
    w = rng.uniform(0, 1, (tokens, used)).astype(np.float32)
    assert np.array_equal(oracle.moe_combine(e, w), reference.moe_combine(e, w))


@pytest.mark.parametrize("wt", H.G8_TYPES)
def test_g8_types_oracle_vs_reference(oracle, reference, wt):
    """Q4_1 / Q5_0 / Q5_K / IQ4_XS on fresh inputs. IQ4_XS: inside the reference kernels' exact domain (see
    tests/golden/make_golden_g8.py: they saturate int16 lanes on large activation quants)."""
    N, K = 33, 1024
    W = H.make_weights(wt, N, K, seed=wt)
    assert np.array_equal(oracle.dequantize(wt, W, N * K), reference.dequantize(wt, W, N * K))
    for T in (1, 5):
        X = H.make_acts(T, K, seed=T)
        if wt == H.IQ4_XS:
            X[:, ::256] = 60.0
        a, b = oracle.mul_mat(W, wt, X), reference.mul_mat(W, wt, X, nth=2)
        assert H.scaled_err(a, b) < 5e-6
    x = (np.random.default_rng(3).standard_normal((7, 512)) * 2).astype(np.float32)
    assert np.array_equal(oracle.quantize_q8_1(x), reference.quantize_q8_1(x))
