"""GPU (B200): the CUDA path, called through the C ABI, against the CPU oracle and the committed reference
fixtures. Bit-exact for activation quantization and integer block dots; scaled error <= 1e-3 (+ NMSE <= 5e-4)
for fp32 outputs. Output buffers are NaN-poisoned first (llamafile/sgemm_matmul_test.cpp:53-56)."""
import numpy as np
import pytest

import helpers as H
from oracle import binding as B

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def gpu_mul_mat(tb, W, wt, X, rows, k):
    w = tb.Weights(W, wt, rows, k)
    x = dev(X)
    out = torch.full((X.shape[0], rows), float("nan"), device="cuda", dtype=torch.float32)
    tb.mul_mat(w, x, X.shape[0], out)
    tb.synchronize()
    res = out.cpu().numpy()
    w.free()
    return res


# ------------------------------------------------------------------ (a) activation quantizers: bit-exact
@pytest.mark.parametrize("rows,k", [(1, 32), (1, 4096), (33, 4096), (7, 14336), (512, 4096), (3, 8192)])
def test_quantize_q8_0_bit_exact(tb, oracle, rows, k):
    x = (np.random.default_rng(rows * k).standard_normal((rows, k)) * 3).astype(np.float32)
    y = torch.zeros(rows * k // 32 * 34, dtype=torch.uint8, device="cuda")
    tb.quantize_q8_0(dev(x), rows, k, y)
    tb.synchronize()
    assert np.array_equal(y.cpu().numpy(), oracle.quantize_q8_0(x))


@pytest.mark.parametrize("rows,k", [(1, 256), (1, 8192), (33, 4096), (5, 28672), (512, 8192)])
def test_quantize_q8_K_bit_exact(tb, oracle, rows, k):
    x = (np.random.default_rng(rows + k).standard_normal((rows, k)) * 0.7).astype(np.float32)
    y = torch.zeros(rows * k // 256 * 292, dtype=torch.uint8, device="cuda")
    tb.quantize_q8_K(dev(x), rows, k, y)
    tb.synchronize()
    assert np.array_equal(y.cpu().numpy(), oracle.quantize_q8_K(x))


def test_quantizers_adversarial_and_golden(tb, oracle, golden):
    for x, g0, gK in ((H.adversarial_rows(256), None, None), (golden["q_adv_x"], golden["q_adv_q8_0"], golden["q_adv_q8_K"]),
                      (golden["q_rnd_x"], golden["q_rnd_q8_0"], golden["q_rnd_q8_K"])):
        rows, k = x.shape
        y0 = torch.zeros(rows * k // 32 * 34, dtype=torch.uint8, device="cuda")
        yK = torch.zeros(rows * k // 256 * 292, dtype=torch.uint8, device="cuda")
        tb.quantize_q8_0(dev(x), rows, k, y0)
        tb.quantize_q8_K(dev(x), rows, k, yK)
        tb.synchronize()
        assert np.array_equal(y0.cpu().numpy(), oracle.quantize_q8_0(x) if g0 is None else g0)
        assert np.array_equal(yK.cpu().numpy(), oracle.quantize_q8_K(x) if gK is None else gK)


def test_quantize_strided_unaligned_and_empty(tb, oracle):
    x = np.random.default_rng(3).standard_normal((5, 301)).astype(np.float32)   # ldx = 301: rows not 16-byte aligned -> scalar loads
    xd = dev(x)
    y = torch.zeros(5 * 8 * 34, dtype=torch.uint8, device="cuda")
    tb.quantize_q8_0(xd, 5, 256, y, ldx=301)
    yk = torch.zeros(5 * 292, dtype=torch.uint8, device="cuda")
    tb.quantize_q8_K(xd, 5, 256, yk, ldx=301)
    tb.quantize_q8_0(xd, 0, 256, y)   # empty input: no-op
    tb.synchronize()
    assert np.array_equal(y.cpu().numpy(), oracle.quantize_q8_0(x[:, :256]))
    assert np.array_equal(yk.cpu().numpy(), oracle.quantize_q8_K(x[:, :256]))


def test_host_pointer_quantizers(tb, oracle, golden):
    x = golden["q_rnd_x"][0]
    assert np.array_equal(tb.quantize_row_q8_0(x), oracle.quantize_q8_0(x[None]))
    assert np.array_equal(tb.quantize_row_q8_K(x), oracle.quantize_q8_K(x[None]))


# ------------------------------------------------------------------ integer block dots: exact
@pytest.mark.parametrize("wt", [H.Q4_0, H.Q8_0, H.Q4_K, H.Q6_K])
def test_block_dots_exact(tb, oracle, wt):
    rows, k = 37, 2048
    W = H.make_weights(wt, rows, k, seed=5, realistic_q4k=False)
    x = H.make_acts(1, k, seed=6)
    Y = oracle.from_float(x, B.VEC_DOT_TYPE[wt])
    nb = k // B.BLCK[wt]
    per = 2 if wt == H.Q4_K else 1
    w = tb.Weights(W, wt, rows, k)
    out = torch.zeros(rows * nb * per, dtype=torch.int32, device="cuda")
    tb.blockdot(w, dev(Y), B.VEC_DOT_TYPE[wt], out)
    tb.synchronize()
    rb = B.row_bytes(wt, k)
    want = np.concatenate([oracle.blockdot(wt, W[i * rb:(i + 1) * rb], Y, nb) for i in range(rows)])
    assert np.array_equal(out.cpu().numpy(), want)


# ------------------------------------------------------------------ (b) decode path vs oracle
@pytest.mark.parametrize("wt", H.WEIGHT_TYPES)
@pytest.mark.parametrize("T", [1, 2, 3, 4, 5, 6, 7, 8])
@pytest.mark.parametrize("N,K", [(64, 256), (131, 1024), (1024, 4096)])
def test_decode_parity(tb, oracle, wt, T, N, K):
    W = H.make_weights(wt, N, K, seed=N + T, realistic_q4k=False)
    X = H.make_acts(T, K, seed=T)
    H.assert_close(gpu_mul_mat(tb, W, wt, X, N, K), oracle.mul_mat(W, wt, X), f"{H.TYPE_NAME[wt]} N={N} K={K} T={T}")


@pytest.mark.parametrize("wt", H.WEIGHT_TYPES)
@pytest.mark.parametrize("T", [1, 2, 3, 8, 9, 33])
def test_golden_mul_mat(tb, golden, wt, T):
    out = gpu_mul_mat(tb, golden[f"w_{wt}"], wt, golden[f"x_{T}"], 48, 512)
    H.assert_close(out, golden[f"mm_{wt}_{T}"], f"golden {H.TYPE_NAME[wt]} T={T}")


@pytest.mark.parametrize("shape", [(64, 64), (256, 64), (64, 256)])
@pytest.mark.parametrize("wt", [H.F16, H.Q8_0])
def test_golden_tinyllama_shapes(tb, golden, wt, shape):
    n, k = shape
    X = golden[f"tl_x_{n}_{k}"]
    H.assert_close(gpu_mul_mat(tb, golden[f"tl_w_{wt}_{n}_{k}"], wt, X, n, k), golden[f"tl_mm_{wt}_{n}_{k}"])
    H.assert_close(gpu_mul_mat(tb, golden[f"tl_w_{wt}_{n}_{k}"], wt, X[:1], n, k), golden[f"tl_mv_{wt}_{n}_{k}"])


@pytest.mark.parametrize("wt,K", [(H.Q4_0, 14336), (H.Q8_0, 14336), (H.Q4_K, 28672), (H.Q6_K, 28672), (H.F16, 14336)])
def test_decode_long_k_chunked_staging(tb, oracle, wt, K):
    """T*K large enough that the activation staging walks K in chunks."""
    N, T = 96, 8
    W = H.make_weights(wt, N, K, seed=2, realistic_q4k=False)
    X = H.make_acts(T, K, seed=3)
    H.assert_close(gpu_mul_mat(tb, W, wt, X, N, K), oracle.mul_mat(W, wt, X))


def test_llama3_full_size_decode_properties(tb, oracle):
    """BASELINE config 2 at full size (N=128256, K=4096, Q4_0, T=1): sampled rows against the oracle, plus the
    size-independent row-slice property: any row range computed alone equals the same rows of the full product
    bit for bit (what row sharding relies on)."""
    N, K = 128256, 4096
    rng = np.random.default_rng(0)
    W = np.empty(N * K // 32 * 18, np.uint8)
    blocks = W.reshape(-1, 18)
    blocks[:, 2:] = rng.integers(0, 256, (blocks.shape[0], 16), dtype=np.uint8)
    blocks[:, :2] = (rng.uniform(0.5, 1.5, blocks.shape[0]) * 2e-3).astype(np.float16).view(np.uint8).reshape(-1, 2)
    X = H.make_acts(1, K)
    full = gpu_mul_mat(tb, W, H.Q4_0, X, N, K)
    rb = K // 32 * 18
    for lo, hi in ((0, 64), (70000, 70032), (N - 40, N)):
        ref = oracle.mul_mat(W[lo * rb:hi * rb], H.Q4_0, X)
        H.assert_close(full[:, lo:hi], ref, f"rows {lo}:{hi}")
        part = gpu_mul_mat(tb, W[lo * rb:hi * rb], H.Q4_0, X, hi - lo, K)
        assert np.array_equal(part, full[:, lo:hi])


# ------------------------------------------------------------------ reference entry points on host pointers
@pytest.mark.parametrize("wt", H.WEIGHT_TYPES)
@pytest.mark.parametrize("n", [1, 4, 8])
def test_llamafile_sgemm_host(tb, oracle, wt, n):
    m, K = 200, 1024
    W = H.make_weights(wt, m, K, seed=n, realistic_q4k=False)
    X = H.make_acts(n, K, seed=10 + n)
    kb = K // B.BLCK[wt]
    vdt = B.VEC_DOT_TYPE[wt]
    ldc = m + 8
    # first attempt with F32 activations: serviced only for F16 weights and n <= 2, exactly like the reference
    C0 = np.full((n, ldc), np.nan, np.float32)
    ok0 = tb.llamafile_sgemm(m, n, kb, W, kb, X, K, C0, ldc, 0, 1, wt, H.F32, H.F32)
    ok_ref, C_ref0 = oracle.sgemm(m, n, kb, W, kb, X, K, ldc, wt, H.F32)
    assert ok0 == ok_ref
    if not ok0:
        assert np.isnan(C0).all()   # returning false must not have touched C
    Bq = oracle.from_float(X, vdt)
    Cm = np.full((n, ldc), np.nan, np.float32)
    assert tb.llamafile_sgemm(m, n, kb, W, kb, Bq, kb, Cm, ldc, 0, 1, wt, vdt, H.F32)
    ok, Cr = oracle.sgemm(m, n, kb, W, kb, Bq, kb, ldc, wt, vdt)
    assert ok
    assert np.isnan(Cm[:, m:]).all()            # padding columns beyond m untouched
    H.assert_close(Cm[:, :m], Cr[:, :m])
    # worker threads other than 0 return true without doing anything (contract in the header)
    C1 = np.full((n, ldc), np.nan, np.float32)
    assert tb.llamafile_sgemm(m, n, kb, W, kb, Bq, kb, C1, ldc, 1, 4, wt, vdt, H.F32)
    assert np.isnan(C1).all()
    tb.weights_cache_clear()


@pytest.mark.parametrize("wt", H.WEIGHT_TYPES)
def test_llamafile_sgemm_host_pinned_buffers(tb, oracle, wt):
    """Decode lane with the caller's buffers already pinned: the kernels read B and write C in place (ldc honoured),
    and repeated calls keep returning complete results (the completion word is a sequence number)."""
    import torch
    m, K, n = 328, 2048, 2
    W = H.make_weights(wt, m, K, seed=3, realistic_q4k=False)
    kb = K // B.BLCK[wt]
    vdt = B.VEC_DOT_TYPE[wt]
    ldc = m + 24
    for rep in range(4):
        X = H.make_acts(n, K, seed=40 + rep)
        Bq = oracle.from_float(X, vdt)
        hB = torch.from_numpy(Bq.view(np.uint8).reshape(-1).copy()).pin_memory()
        hC = torch.full((n, ldc), float("nan")).pin_memory()
        assert tb.llamafile_sgemm(m, n, kb, W, kb, hB.numpy(), kb, hC.numpy(), ldc, 0, 1, wt, vdt, H.F32)
        ok, Cr = oracle.sgemm(m, n, kb, W, kb, Bq, kb, ldc, wt, vdt)
        assert ok
        Cm = hC.numpy()
        assert np.isnan(Cm[:, m:]).all()
        H.assert_close(Cm[:, :m], Cr[:, :m], f"pinned rep {rep}")
    tb.weights_cache_clear()


def test_llamafile_sgemm_rejects_like_reference(tb):
    W = np.zeros(18 * 8 * 4, np.uint8); X = np.zeros((1, 256), np.float32); Cm = np.full((1, 4), np.nan, np.float32)
    assert not tb.llamafile_sgemm(4, 1, 8, W, 8, X, 256, Cm, 4, 0, 1, H.Q4_0, H.F32, H.F32)    # WANT_QUANTIZATION
    assert not tb.llamafile_sgemm(4, 1, 8, W, 8, X, 8, Cm, 4, 0, 1, H.Q4_0, H.Q8_0, H.F16)     # Ctype
    assert not tb.llamafile_sgemm(4, 1, 8, W, 4, X, 8, Cm, 4, 0, 1, H.Q4_0, H.Q8_0, H.F32)     # lda < k
    assert np.isnan(Cm).all()


# ------------------------------------------------------------------ (d) mixmul
def run_mixmul_host(tb, W, wt, X, plan, rows, cols, experts, fill=np.nan):
    tokens, tasks, _ = X.shape
    thinkers = plan.shape[1]
    out = np.full((tokens, thinkers, rows), fill, np.float32)
    tw, tx = tb.GgmlTensor.of(W, wt, [cols, rows, experts]), tb.GgmlTensor.of(X, H.F32, [cols, tasks, tokens])
    tp, to = tb.GgmlTensor.of(plan, H.I32, [thinkers, tokens]), tb.GgmlTensor.of(out, H.F32, [rows, thinkers, tokens])
    need = tb.llamafile_mixmul_needs(tw, tx, tp)
    scratch = np.zeros(need + 4096, np.uint8)
    params = tb.GgmlComputeParams(0, 1, scratch.size, scratch.ctypes.data, None)
    ok = tb.llamafile_mixmul(params, tw, tx, tp, to)
    tb.weights_cache_clear()
    return ok, out


@pytest.mark.parametrize("wt", [H.Q4_0, H.Q8_0, H.F16])
@pytest.mark.parametrize("tokens,tasks", [(1, 1), (5, 1), (5, 2), (19, 2)])
def test_golden_mixmul(tb, golden, wt, tokens, tasks):
    ok, out = run_mixmul_host(tb, golden[f"mx_w_{wt}"], wt, golden[f"mx_x_{tokens}_{tasks}"], golden[f"mx_plan_{tokens}"], 32, 256, 4)
    assert ok
    H.assert_close(out, golden[f"mx_out_{wt}_{tokens}_{tasks}"])


@pytest.mark.parametrize("wt", [H.Q4_0, H.Q8_0, H.F16])
@pytest.mark.parametrize("tokens,tasks,skew", [(1, 1, False), (8, 1, False), (8, 2, False), (64, 1, False), (40, 2, True)])
def test_mixmul_vs_oracle(tb, oracle, wt, tokens, tasks, skew):
    rows, cols, experts, thinkers = 96, 1024, 8, 2
    W = H.make_weights(wt, experts * rows, cols, seed=tokens)
    X = np.random.default_rng(tokens).standard_normal((tokens, tasks, cols)).astype(np.float32)
    plan = H.make_plan(tokens, experts, thinkers, skew=skew)
    ok, out = run_mixmul_host(tb, W, wt, X, plan, rows, cols, experts)
    assert ok
    ref = np.full((tokens, thinkers, rows), np.nan, np.float32)
    assert oracle.mixmul(B.Tensor.of(W, wt, [cols, rows, experts]), B.Tensor.of(X, H.F32, [cols, tasks, tokens]),
                         B.Tensor.of(plan, H.I32, [thinkers, tokens]), B.Tensor.of(ref, H.F32, [rows, thinkers, tokens]))
    H.assert_close(out, ref)


def test_mixmul_invalid_expert_rows_untouched_and_unsupported_types(tb, oracle):
    rows, cols, experts = 16, 256, 4
    W = H.make_weights(H.Q8_0, experts * rows, cols)
    X = np.random.default_rng(1).standard_normal((3, 1, cols)).astype(np.float32)
    plan = np.array([[0, 1], [9, 2], [3, -1]], np.int32)
    ok, out = run_mixmul_host(tb, W, H.Q8_0, X, plan, rows, cols, experts, fill=7.5)
    assert ok and (out[1, 0] == 7.5).all() and (out[2, 1] == 7.5).all()
    ref = np.full((3, 2, rows), 7.5, np.float32)
    assert oracle.mixmul(B.Tensor.of(W, H.Q8_0, [cols, rows, experts]), B.Tensor.of(X, H.F32, [cols, 1, 3]),
                         B.Tensor.of(plan, H.I32, [2, 3]), B.Tensor.of(ref, H.F32, [rows, 2, 3]))
    H.assert_close(out, ref)
    # Q4_K experts are not llamafile_mixmul's job (they go through llamafile_mixmul_iqk)
    ok, _ = run_mixmul_host(tb, H.synth_q4_K(experts * rows, cols, 1), H.Q4_K, X, np.zeros((3, 2), np.int32), rows, cols, experts)
    assert not ok


def test_mixmul_device_api_no_host_sync(tb, oracle):
    """tb200_mixmul_dev with the routing plan resident on the device (Mixtral-like widths, decode and small batch)."""
    rows, cols, experts, thinkers = 512, 4096, 8, 2
    W = H.make_weights(H.Q4_0, experts * rows, cols, seed=4)
    w = tb.Weights(W, H.Q4_0, experts * rows, cols)
    for tokens in (1, 8, 64):
        X = np.random.default_rng(tokens).standard_normal((tokens, 1, cols)).astype(np.float32)
        plan = H.make_plan(tokens, experts, thinkers, seed=tokens)
        out = torch.full((tokens, thinkers, rows), float("nan"), device="cuda")
        assert tb.mixmul(w, experts, dev(X), 1, dev(plan), thinkers, tokens, out)
        tb.synchronize()
        ref = np.full((tokens, thinkers, rows), np.nan, np.float32)
        assert oracle.mixmul(B.Tensor.of(W, H.Q4_0, [cols, rows, experts]), B.Tensor.of(X, H.F32, [cols, 1, tokens]),
                             B.Tensor.of(plan, H.I32, [thinkers, tokens]), B.Tensor.of(ref, H.F32, [rows, thinkers, tokens]))
        H.assert_close(out.cpu().numpy(), ref, f"tokens={tokens}")
    w.free()


@pytest.mark.parametrize("tokens,skew", [(5, False), (12, False), (24, True), (60, False)])
def test_mixmul_grouped_resident_odd_rows(tb, oracle, tokens, skew):
    """Q4_0 decode-size routing through the grouped resident kernel (items of up to 8 tokens, CTAs spread over the live items):
    odd row count per expert (row-pair tail), experts with more than 8 tokens (several items), experts with none."""
    rows, cols, experts, thinkers = 97, 2048, 8, 2
    W = H.make_weights(H.Q4_0, experts * rows, cols, seed=tokens + 1)
    w = tb.Weights(W, H.Q4_0, experts * rows, cols)
    X = np.random.default_rng(tokens).standard_normal((tokens, 1, cols)).astype(np.float32)
    plan = H.make_plan(tokens, experts, thinkers, skew=skew, seed=tokens)
    out = torch.full((tokens, thinkers, rows), float("nan"), device="cuda")
    assert tb.mixmul(w, experts, dev(X), 1, dev(plan), thinkers, tokens, out)
    tb.synchronize()
    ref = np.full((tokens, thinkers, rows), np.nan, np.float32)
    assert oracle.mixmul(B.Tensor.of(W, H.Q4_0, [cols, rows, experts]), B.Tensor.of(X, H.F32, [cols, 1, tokens]),
                         B.Tensor.of(plan, H.I32, [thinkers, tokens]), B.Tensor.of(ref, H.F32, [rows, thinkers, tokens]))
    H.assert_close(out.cpu().numpy(), ref, f"tokens={tokens}")
    w.free()


@pytest.mark.parametrize("typeA", [H.Q4_K, H.Q4_0, H.Q6_K])
def test_llamafile_mixmul_iqk_host(tb, oracle, golden, typeA):
    if typeA == H.Q4_K:
        W, X, mapping = golden["iqk_w"], golden["iqk_x"], golden["iqk_map"]
        Nx, ne00, ne11, ne12 = 40, 512, 2, 6
    else:
        Nx, ne00, ne11, ne12 = 24, 512, 2, 9
        W = H.make_weights(typeA, Nx, ne00, seed=3)
        X = np.random.default_rng(4).standard_normal((ne12 * ne11, ne00)).astype(np.float32)
        mapping = np.array([[i % 2, (3 * i) % ne12] for i in range(11)], np.int32)
    Bq = oracle.from_float(X, B.VEC_DOT_TYPE[typeA])
    out = np.full((ne12, 2, Nx), np.nan, np.float32)
    assert tb.llamafile_mixmul_iqk(Nx, len(mapping), ne00, ne11, typeA, W, Bq, out, Nx * 4, 2 * Nx * 4, mapping, 0, 1)
    ref = np.full((ne12, 2, Nx), np.nan, np.float32)
    assert oracle.mixmul_iqk(Nx, len(mapping), ne00, ne11, typeA, W, Bq, ref, Nx * 4, 2 * Nx * 4, mapping)
    if typeA == H.Q4_K:
        ref = golden["iqk_out"]
    assert np.array_equal(np.isnan(out), np.isnan(ref))
    m = ~np.isnan(ref)
    H.assert_close(out[m], ref[m])
    tb.weights_cache_clear()


# ------------------------------------------------------------------ (c) prefill path (tcgen05) vs oracle
@pytest.mark.parametrize("wt", H.WEIGHT_TYPES)
@pytest.mark.parametrize("N,K,T", [(128, 256, 32), (256, 1024, 64), (130, 512, 33), (384, 4096, 512), (1000, 2048, 300), (256, 512, 9)])
def test_prefill_parity(tb, oracle, wt, N, K, T):
    W = H.make_weights(wt, N, K, seed=N + T, realistic_q4k=False)
    X = H.make_acts(T, K, seed=T)
    H.assert_close(gpu_mul_mat(tb, W, wt, X, N, K), oracle.mul_mat(W, wt, X), f"prefill {H.TYPE_NAME[wt]} N={N} K={K} T={T}")


_PAIR_SNIPPET = r"""
import sys
sys.path.insert(0, {root!r}); sys.path.insert(0, {root!r} + "/tests")
import numpy as np, torch
import helpers as H
import llamafile_b200 as tb
from oracle import binding as B
assert tb.link()
o = B.oracle()
for wt in H.WEIGHT_TYPES:
    for (N, K, T) in ((384, 1024, 300), (1000, 512, 129)):
        W = H.make_weights(wt, N, K, seed=N + T, realistic_q4k=False)
        X = H.make_acts(T, K, seed=T)
        w = tb.Weights(W, wt, N, K)
        out = torch.full((T, N), float("nan"), device="cuda")
        tb.mul_mat(w, torch.from_numpy(X).cuda(), T, out)
        tb.synchronize()
        H.assert_close(out.cpu().numpy(), o.mul_mat(W, wt, X), f"pair kernel {{H.TYPE_NAME[wt]}} N={{N}} K={{K}} T={{T}}")
print("PAIR_OK")
"""


def test_prefill_cta_pair_kernel_all_types():
    """The cta_group::2 kernel is the default only for F16 weights; TB200_GEMM_PAIR=2 (read once per process) routes every
    weight type through it, including an odd number of row tiles (the pair's second CTA is all padding)."""
    import os
    import subprocess
    import sys
    root = str(__import__("pathlib").Path(__file__).resolve().parent.parent)
    env = dict(os.environ, TB200_GEMM_PAIR="2")
    r = subprocess.run([sys.executable, "-c", _PAIR_SNIPPET.format(root=root)], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "PAIR_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


_MMA_SNIPPET = r"""
import sys
sys.path.insert(0, {root!r}); sys.path.insert(0, {root!r} + "/tests")
import numpy as np, torch
import helpers as H
import llamafile_b200 as tb
from oracle import binding as B
assert tb.link()
o = B.oracle()
for (N, K) in ((131, 1024), (1024, 4096), (50, 14336)):
    W = H.make_weights(H.Q4_0, N, K, seed=N)
    w = tb.Weights(W, H.Q4_0, N, K)
    for T in (4, 5, 8):
        X = H.make_acts(T, K, seed=T)
        out = torch.full((T, N), float("nan"), device="cuda")
        tb.mul_mat(w, torch.from_numpy(X).cuda(), T, out)
        tb.synchronize()
        H.assert_close(out.cpu().numpy(), o.mul_mat(W, H.Q4_0, X), f"mma N={{N}} K={{K}} T={{T}}")
print("MMA_OK")
"""


@pytest.mark.parametrize("split", ["0", "1"])
def test_decode_mma_kernel_both_forms(split):
    """Q4_0 at 4..8 tokens runs on mma.sync m16n8k32: TB200_MMA_SPLIT (read once per process) forces the warp-per-tile form (0,
    otherwise only used for very tall matrices) or the split-K form (1); ragged row tiles, K slices shorter than a batch."""
    import os
    import subprocess
    import sys
    root = str(__import__("pathlib").Path(__file__).resolve().parent.parent)
    env = dict(os.environ, TB200_MMA_SPLIT=split)
    r = subprocess.run([sys.executable, "-c", _MMA_SNIPPET.format(root=root)], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "MMA_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_decode_mma_exact_integer_path(tb, oracle):
    """Activations that quantize exactly and power-of-two scales: every fp32 product and sum of the tensor-core path is
    exact, so the result must EQUAL the oracle's (the int32 block sums and the -8 correction carry no rounding)."""
    N, K, T = 64, 1024, 8
    rng = np.random.default_rng(11)
    blocks = np.zeros((N * K // 32, 18), np.uint8)
    blocks[:, :2] = np.full(N * K // 32, 2.0 ** -6, np.float16).view(np.uint8).reshape(-1, 2)
    blocks[:, 2:] = rng.integers(0, 256, (N * K // 32, 16), dtype=np.uint8)
    W = blocks.reshape(-1)
    # every 32-block holds +-127 once, so d = amax/127 = 1 and the int8 quants are the values themselves
    X = rng.integers(-20, 21, (T, K)).astype(np.float32)
    X[:, ::32] = 127.0
    out = gpu_mul_mat(tb, W, H.Q4_0, X, N, K)
    ref = oracle.mul_mat(W, H.Q4_0, X)
    assert np.array_equal(out, ref), float(np.abs(out - ref).max())


def test_prefill_activation_outliers(tb, oracle):
    """Token rows with huge / tiny magnitudes: the per-token power-of-two scaling keeps fp16 operands in range."""
    N, K, T = 256, 1024, 48
    W = H.make_weights(H.Q4_0, N, K, seed=1)
    X = H.make_acts(T, K, seed=2)
    X[3] *= 3.0e5
    X[7] *= 1.0e-6
    X[11, 5] = 9.0e4
    out, ref = gpu_mul_mat(tb, W, H.Q4_0, X, N, K), oracle.mul_mat(W, H.Q4_0, X)
    for t in range(T):   # per token, since magnitudes differ by 11 orders
        H.assert_close(out[t], ref[t], f"token {t}")


def test_prefill_llama_shape_sampled(tb, oracle):
    """Llama-3-8B prefill shape at full K and N: every token row, a sample of output columns."""
    N, K, T = 4096, 4096, 512
    W = H.make_weights(H.Q4_0, N, K, seed=7)
    X = H.make_acts(T, K, seed=8)
    out = gpu_mul_mat(tb, W, H.Q4_0, X, N, K)
    rb = K // 32 * 18
    for lo, hi in ((0, 128), (2000, 2100), (N - 64, N)):
        H.assert_close(out[:, lo:hi], oracle.mul_mat(W[lo * rb:hi * rb], H.Q4_0, X), f"cols {lo}:{hi}")


# ------------------------------------------------------------------ decode API variants
@pytest.mark.parametrize("wt", H.WEIGHT_TYPES)
@pytest.mark.parametrize("T", [1, 3, 8])
def test_resident_activations_api(tb, oracle, wt, T):
    """tb200_acts_quantize + tb200_mul_mat_acts (planar activations) == one-shot tb200_mul_mat == oracle."""
    K = 1024
    X = H.make_acts(T, K, seed=T)
    acts = tb.Acts(wt, K, 8)
    xd = dev(X)
    acts.quantize(xd, T)
    for N in (96, 320):
        W = H.make_weights(wt, N, K, seed=N, realistic_q4k=False)
        w = tb.Weights(W, wt, N, K)
        out = torch.full((T, N), float("nan"), device="cuda")
        tb.mul_mat_acts(w, acts, out)
        tb.synchronize()
        H.assert_close(out.cpu().numpy(), oracle.mul_mat(W, wt, X))
        w.free()
    acts.free()


@pytest.mark.parametrize("wt", [H.Q4_0, H.Q4_K, H.Q6_K])
@pytest.mark.parametrize("T", [1, 2, 8])
def test_mul_mat_multi(tb, oracle, wt, T):
    """Several matrices sharing K and the activations in one launch (QKV / gate-up); uneven row counts."""
    K = 2048
    X = H.make_acts(T, K, seed=T + 40)
    shapes = [520, 128, 130]
    Ws = [H.make_weights(wt, n, K, seed=n, realistic_q4k=False) for n in shapes]
    ws = [tb.Weights(W, wt, n, K) for W, n in zip(Ws, shapes)]
    outs = [torch.full((T, n), float("nan"), device="cuda") for n in shapes]
    tb.mul_mat_multi(ws, dev(X), T, outs)
    tb.synchronize()
    for W, n, o in zip(Ws, shapes, outs):
        H.assert_close(o.cpu().numpy(), oracle.mul_mat(W, wt, X), f"multi n={n}")
    for w in ws:
        w.free()


# ------------------------------------------------------------------ mixmul at prefill sizes (grouped tcgen05 launch)
@pytest.mark.parametrize("wt", [H.Q4_0, H.Q8_0, H.F16])
@pytest.mark.parametrize("tokens,tasks,skew", [(100, 1, False), (300, 2, False), (256, 1, True)])
def test_mixmul_prefill_vs_oracle(tb, oracle, wt, tokens, tasks, skew):
    rows, cols, experts, thinkers = 200, 1024, 8, 2
    W = H.make_weights(wt, experts * rows, cols, seed=tokens)
    X = np.random.default_rng(tokens).standard_normal((tokens, tasks, cols)).astype(np.float32)
    plan = H.make_plan(tokens, experts, thinkers, skew=skew)
    ok, out = run_mixmul_host(tb, W, wt, X, plan, rows, cols, experts)
    assert ok
    ref = np.full((tokens, thinkers, rows), np.nan, np.float32)
    assert oracle.mixmul(B.Tensor.of(W, wt, [cols, rows, experts]), B.Tensor.of(X, H.F32, [cols, tasks, tokens]),
                         B.Tensor.of(plan, H.I32, [thinkers, tokens]), B.Tensor.of(ref, H.F32, [rows, thinkers, tokens]))
    H.assert_close(out, ref)


def test_mixmul_prefill_invalid_experts_untouched(tb, oracle):
    rows, cols, experts, thinkers, tokens = 128, 512, 4, 2, 200
    W = H.make_weights(H.Q4_0, experts * rows, cols, seed=2)
    X = np.random.default_rng(5).standard_normal((tokens, 1, cols)).astype(np.float32)
    plan = H.make_plan(tokens, experts, thinkers, seed=6)
    plan[::7, 0] = 11
    plan[3::9, 1] = -2
    ok, out = run_mixmul_host(tb, W, H.Q4_0, X, plan, rows, cols, experts, fill=7.5)
    assert ok
    ref = np.full((tokens, thinkers, rows), 7.5, np.float32)
    assert oracle.mixmul(B.Tensor.of(W, H.Q4_0, [cols, rows, experts]), B.Tensor.of(X, H.F32, [cols, 1, tokens]),
                         B.Tensor.of(plan, H.I32, [thinkers, tokens]), B.Tensor.of(ref, H.F32, [rows, thinkers, tokens]))
    assert np.array_equal(out == 7.5, ref == 7.5)
    H.assert_close(out, ref)
