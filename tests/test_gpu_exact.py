"""GPU (B200): exact-arithmetic parity of the HOT decode kernels, and sampled parity at BASELINE.json's full sizes.

north_star asks for bit-exact integer block dot products (ggml-quants.c:4192-4206 Q4_0, :5227 Q8_0, :7341-7397 Q4_K,
:8642-8682 Q6_K). A separate checker kernel cannot prove that for the kernels that run, so every case below builds
inputs on which every fp32 operation of the path is exact — power-of-two weight scales, activations that quantize to
themselves (each block holds +-127 once, so d_y = 1 and q = x) — and demands np.array_equal with the oracle: any
slip in the -8 fold, the `sh >> 4` of the in-place high nibbles, the packed 6-bit scale decode or the 4*(q-32) form of
Q6_K changes an integer and therefore a bit of the result. The shapes are chosen so that each launcher branch of
tb200_launch_gemv is taken (tb200_gemv.cu launch_T)."""
import os
import subprocess
import sys
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


def exact_acts(T, K, seed, block, lim=12):
    """Integers in [-lim, lim] with one 127 per `block` elements: Q8_0 gives d = 1, q = x; Q8_K gives iscale = -1, d = -1, q = -x."""
    rng = np.random.default_rng(seed)
    X = rng.integers(-lim, lim + 1, (T, K)).astype(np.float32)
    X[:, ::block] = 127.0
    return X


def exact_weights(wt, N, K, seed):
    """Random quant bytes, every floating-point scale a power of two, sub-block scales any integer the format holds."""
    rng = np.random.default_rng(seed)
    if wt == H.Q4_0:
        b = rng.integers(0, 256, (N * K // 32, 18), dtype=np.uint8)
        b[:, :2] = np.full(len(b), 2.0 ** -6, np.float16).view(np.uint8).reshape(-1, 2)
    elif wt == H.Q8_0:
        b = rng.integers(0, 256, (N * K // 32, 34), dtype=np.uint8)
        b[:, :2] = np.full(len(b), 2.0 ** -7, np.float16).view(np.uint8).reshape(-1, 2)
    elif wt == H.Q4_K:
        b = rng.integers(0, 256, (N * K // 256, 144), dtype=np.uint8)
        b[:, 0:2] = np.full(len(b), 2.0 ** -8, np.float16).view(np.uint8).reshape(-1, 2)
        b[:, 2:4] = np.full(len(b), 2.0 ** -7, np.float16).view(np.uint8).reshape(-1, 2)
    elif wt == H.Q6_K:
        b = rng.integers(0, 256, (N * K // 256, 210), dtype=np.uint8)
        b[:, 208:210] = np.full(len(b), 2.0 ** -10, np.float16).view(np.uint8).reshape(-1, 2)
    else:
        raise ValueError(wt)
    return b.reshape(-1)


# (weight type, N, K, T) -> the kernel that serves it (launch_T in tb200_gemv.cu)
EXACT_CASES = [
    (H.Q4_0, 64, 1024, 1, "gemv_q4_0_kernel<1,1>"), (H.Q4_0, 64, 1024, 2, "gemv_q4_0_kernel<2,1>"), (H.Q4_0, 131, 1024, 3, "gemv_q4_0_kernel<3,1>"),
    (H.Q4_0, 5000, 1024, 1, "gemv_q4_0_kernel<1,2>"), (H.Q4_0, 5001, 1024, 3, "gemv_q4_0_kernel<3,2>"),
    (H.Q4_0, 30001, 1024, 1, "gemv_q4_0_resident_kernel<1>"), (H.Q4_0, 30000, 1024, 2, "gemv_q4_0_resident_kernel<2>"),
    (H.Q4_0, 200, 1024, 4, "gemm_q4_0_mma_kernel<4>"), (H.Q4_0, 200, 2048, 7, "gemm_q4_0_mma_kernel<7>"),
    (H.Q8_0, 64, 1024, 1, "gemv_q8_0_kernel<1,1>"), (H.Q8_0, 5000, 512, 1, "gemv_q8_0_kernel<1,2>"),
    (H.Q8_0, 130, 1024, 2, "gemm_q8_0_mma_kernel<2>"), (H.Q8_0, 130, 1024, 5, "gemm_q8_0_mma_kernel<5>"), (H.Q8_0, 4100, 1024, 8, "gemm_q8_0_mma_kernel<8>"),
    (H.Q4_K, 64, 1024, 1, "gemv_q4_K_kernel<1,1>"), (H.Q4_K, 5000, 1024, 1, "gemv_q4_K_kernel<1,2>"), (H.Q4_K, 131, 2048, 3, "gemv_q4_K_kernel<3,1>"),
    (H.Q4_K, 96, 1024, 8, "gemv_q4_K_kernel<8,1>"),
    (H.Q6_K, 64, 1024, 1, "gemv_q6_K_kernel<1,1>"), (H.Q6_K, 5000, 1024, 2, "gemv_q6_K_kernel<2,2>"), (H.Q6_K, 131, 2048, 3, "gemv_q6_K_kernel<3,1>"),
    (H.Q6_K, 96, 1024, 8, "gemv_q6_K_kernel<8,1>"),
]


@pytest.mark.parametrize("wt,N,K,T,kernel", EXACT_CASES, ids=[c[4] + f"-N{c[1]}" for c in EXACT_CASES])
def test_hot_kernel_exact_integer_arithmetic(tb, oracle, wt, N, K, T, kernel):
    W = exact_weights(wt, N, K, seed=N + T)
    X = exact_acts(T, K, seed=T, block=B.BLCK[wt], lim=6 if wt == H.Q6_K else 12)
    out = gpu_mul_mat(tb, W, wt, X, N, K)
    ref = oracle.mul_mat(W, wt, X)
    assert np.array_equal(out, ref), f"{kernel}: max |diff| {float(np.abs(out - ref).max())}"


@pytest.mark.parametrize("wt", [H.Q4_0, H.Q4_K, H.Q6_K])
def test_multi_matrix_launch_exact(tb, oracle, wt):
    """The wq/wk/wv / gate-up form (several matrices per launch) on exact inputs."""
    K, T = 1024, 1
    shapes = [520, 129, 4100]
    Ws = [exact_weights(wt, n, K, seed=n) for n in shapes]
    X = exact_acts(T, K, seed=9, block=B.BLCK[wt], lim=6 if wt == H.Q6_K else 12)
    ws = [tb.Weights(W, wt, n, K) for W, n in zip(Ws, shapes)]
    outs = [torch.full((T, n), float("nan"), device="cuda") for n in shapes]
    tb.mul_mat_multi(ws, dev(X), T, outs)
    tb.synchronize()
    for W, n, o in zip(Ws, shapes, outs):
        assert np.array_equal(o.cpu().numpy(), oracle.mul_mat(W, wt, X)), f"multi n={n}"
    for w in ws:
        w.free()


@pytest.mark.parametrize("tokens", [12, 24])
def test_grouped_mixmul_kernel_exact(tb, oracle, tokens):
    """gemv_q4_0_grouped_kernel (mixmul decode, more routed pairs than experts) on exact inputs."""
    rows, cols, experts, thinkers = 96, 1024, 8, 2
    W = exact_weights(H.Q4_0, experts * rows, cols, seed=tokens)
    X = exact_acts(tokens, cols, seed=3, block=32).reshape(tokens, 1, cols)
    plan = H.make_plan(tokens, experts, thinkers, seed=tokens)
    w = tb.Weights(W, H.Q4_0, experts * rows, cols)
    out = torch.full((tokens, thinkers, rows), float("nan"), device="cuda")
    assert tb.mixmul(w, experts, dev(X), 1, dev(plan), thinkers, tokens, out)
    tb.synchronize()
    ref = np.full((tokens, thinkers, rows), np.nan, np.float32)
    assert oracle.mixmul(B.Tensor.of(W, H.Q4_0, [cols, rows, experts]), B.Tensor.of(X, H.F32, [cols, 1, tokens]),
                         B.Tensor.of(plan, H.I32, [thinkers, tokens]), B.Tensor.of(ref, H.F32, [rows, thinkers, tokens]))
    assert np.array_equal(out.cpu().numpy(), ref)
    w.free()


# ------------------------------------------------------------------ a7: ggml_fp32_to_fp16_row on the GPU, bit for bit
def test_fp32_to_fp16_kernel_bit_exact(tb, oracle, golden):
    x = golden["q_rnd_x"]
    got = np.concatenate([tb.fp32_to_fp16_row(r).view(np.uint8) for r in x])
    assert np.array_equal(got, golden["q_rnd_f16"])
    # rounding corners: ties to even, fp16 subnormals, overflow to inf, signed zeros
    c = np.array([1.0 + 2.0 ** -11, 1.0 + 3 * 2.0 ** -11, 2.0 ** -24, 1.5 * 2.0 ** -24, 2.0 ** -25, 65504.0, 65519.9, 65520.0, 1e9, -0.0, 0.0,
                  -1e-8, 6.1e-5, 6.0e-8, -65520.0, 3.14159265], np.float32)
    assert np.array_equal(tb.fp32_to_fp16_row(c).view(np.uint8), oracle.fp32_to_fp16(c[None]).view(np.uint8).reshape(-1))


# ------------------------------------------------------------------ BASELINE sizes, sampled rows against the oracle
def synth_bytes(wt, N, K, seed):
    """Random quant bytes with realistic scales (fast: no per-element quantizer pass over 10^8 weights)."""
    rng = np.random.default_rng(seed)
    if wt == H.F16:
        return (rng.standard_normal((N, K), dtype=np.float32) * 0.02).astype(np.float16).view(np.uint8).reshape(-1)
    ts, blck = B.TYPE_SIZE[wt], B.BLCK[wt]
    b = rng.integers(0, 256, (N * K // blck, ts), dtype=np.uint8)
    d = (rng.uniform(0.5, 1.5, len(b)) * 2e-3).astype(np.float16).view(np.uint8).reshape(-1, 2)
    if wt == H.Q6_K:
        b[:, 208:210] = (rng.uniform(0.5, 1.5, len(b)) * 2e-5).astype(np.float16).view(np.uint8).reshape(-1, 2)
    else:
        b[:, :2] = d
    if wt == H.Q4_K:
        b[:, 2:4] = (rng.uniform(0.5, 1.5, len(b)) * 4e-3).astype(np.float16).view(np.uint8).reshape(-1, 2)
    return b.reshape(-1)


def check_sampled(tb, oracle, wt, N, K, X, spans, what):
    W = synth_bytes(wt, N, K, seed=N + K)
    out = gpu_mul_mat(tb, W, wt, X, N, K)
    assert not np.isnan(out).any(), f"{what}: unwritten outputs"
    rb = B.row_bytes(wt, K)
    for lo, hi in spans:
        H.assert_close(out[:, lo:hi], oracle.mul_mat(W[lo * rb:hi * rb], wt, X), f"{what} rows {lo}:{hi}")


@pytest.mark.parametrize("wt", [H.Q8_0, H.Q4_0, H.Q4_K, H.F16])
def test_prefill_baseline_size_T4096(tb, oracle, wt):
    """BASELINE config 3 at its largest point (T=4096, N=14336, K=4096): the BN=256, 12-wave configuration."""
    N, K, T = 14336, 4096, 4096
    X = H.make_acts(T, K, seed=5)
    check_sampled(tb, oracle, wt, N, K, X, ((0, 24), (7777, 7801), (N - 24, N)), f"prefill {H.TYPE_NAME[wt]} T=4096")


@pytest.mark.parametrize("N,K", [(8192, 8192), (28672, 8192), (8192, 28672)])
@pytest.mark.parametrize("T", [1, 512])
def test_llama70b_q4_K_shapes(tb, oracle, N, K, T):
    """BASELINE config 5 shapes (Q4_K, Llama-3-70B), decode and prefill, one GPU."""
    X = H.make_acts(T, K, seed=T)
    check_sampled(tb, oracle, H.Q4_K, N, K, X, ((0, 16), (N // 2 + 5, N // 2 + 21), (N - 16, N)), f"70B q4_K N={N} K={K} T={T}")


@pytest.mark.parametrize("tokens", [1, 8, 64, 512])
def test_mixtral_mul_mat_id_baseline_size(tb, oracle, tokens):
    """BASELINE config 4: Mixtral-8x7B ffn_gate_exps (8 experts x 14336 x 4096, Q4_0), top-2, sampled (token, expert, rows)."""
    rows, cols, experts, thinkers = 14336, 4096, 8, 2
    W = synth_bytes(H.Q4_0, experts * rows, cols, seed=77)
    w = tb.Weights(W, H.Q4_0, experts * rows, cols)
    X = np.random.default_rng(tokens).standard_normal((tokens, 1, cols)).astype(np.float32)
    plan = H.make_plan(tokens, experts, thinkers, seed=tokens + 1)
    out = torch.full((tokens, thinkers, rows), float("nan"), device="cuda")
    assert tb.mixmul(w, experts, dev(X), 1, dev(plan), thinkers, tokens, out)
    tb.synchronize()
    got = out.cpu().numpy()
    w.free()
    assert not np.isnan(got).any()
    rb = B.row_bytes(H.Q4_0, cols)
    for t in sorted({0, tokens // 3, tokens - 1}):
        for th in range(thinkers):
            e = int(plan[t, th])
            for lo in (0, 9000, rows - 32):
                ref = oracle.mul_mat(W[(e * rows + lo) * rb:(e * rows + lo + 32) * rb], H.Q4_0, X[t])
                H.assert_close(got[t, th, lo:lo + 32][None], ref, f"tokens={tokens} t={t} expert={e} rows {lo}")


# ------------------------------------------------------------------ prefill: extreme WEIGHT scales (A' = fp16(d * (q - 8)))
@pytest.mark.parametrize("wt", [H.Q4_0, H.Q8_0])
@pytest.mark.parametrize("scale,name", [(2.0 ** -22, "fp16-subnormal products"), (3.0e-6, "near the subnormal edge"), (65504.0 / 8, "d*q at fp16 max"), (1.0, "unit")])
def test_prefill_weight_scale_extremes(tb, oracle, wt, scale, name):
    N, K, T = 256, 1024, 64
    rng = np.random.default_rng(4)
    ts = B.TYPE_SIZE[wt]
    b = rng.integers(0, 256, (N * K // 32, ts), dtype=np.uint8)
    sc = scale / (16 if wt == H.Q8_0 else 1)   # |q| <= 127 for Q8_0: keep d*q inside fp16 for the "max" case
    b[:, :2] = (rng.uniform(0.6, 1.0, len(b)) * sc).astype(np.float16).view(np.uint8).reshape(-1, 2)
    W = b.reshape(-1)
    X = H.make_acts(T, K, seed=6)
    H.assert_close(gpu_mul_mat(tb, W, wt, X, N, K), oracle.mul_mat(W, wt, X), f"{H.TYPE_NAME[wt]} weights, {name}")


def test_prefill_realistic_kquant_weights(tb, oracle, golden):
    """Prefill against reference-quantised K-quant weights (the golden file's w_12 / w_14 came out of quantize_q4_K / q6_K)."""
    for wt in (H.Q4_K, H.Q6_K):
        out = gpu_mul_mat(tb, golden[f"w_{wt}"], wt, golden["x_33"], 48, 512)
        H.assert_close(out, golden[f"mm_{wt}_33"], f"golden prefill {H.TYPE_NAME[wt]}")


# ------------------------------------------------------------------ multi-GPU: fused all-gather parity (skips on one GPU)
@pytest.mark.parametrize("world", [2, 8])
def test_multi_gpu_fused_gather_parity(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    port = 29600 + os.getpid() % 300 + world
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
                        "--master-port", str(port), str(ROOT / "tools" / "check_p2p.py")], capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0 and "P2P_GATHER_OK" in r.stdout and "P2P_PLAN_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


# ------------------------------------------------------------------ host entry point: what may be kept resident
@pytest.mark.parametrize("wt", [H.F16, H.Q8_0, H.Q4_0])
def test_llamafile_sgemm_sees_in_place_updates_of_unregistered_A(tb, oracle, wt):
    """ggml also sends the KV cache (F16, or Q8_0/Q4_0 with -ctk) through llamafile_sgemm as A: the same address and shape
    recur for many tokens with different bytes. Only registered (immutable) ranges may be cached."""
    m, K, n = 96, 512, 1
    kb = K // B.BLCK[wt]
    vdt = B.VEC_DOT_TYPE[wt]
    A = H.make_weights(wt, m, K, seed=1).copy()
    X = H.make_acts(n, K, seed=2)
    Bq = oracle.from_float(X, vdt)
    before = tb.weights_cache_bytes()
    for step in range(3):
        A[:] = H.make_weights(wt, m, K, seed=10 + step)           # same address, new contents
        Cm = np.full((n, m), np.nan, np.float32)
        assert tb.llamafile_sgemm(m, n, kb, A, kb, Bq, kb, Cm, m, 0, 1, wt, vdt, H.F32)
        ok, Cr = oracle.sgemm(m, n, kb, A, kb, Bq, kb, m, wt, vdt)
        assert ok
        H.assert_close(Cm, Cr, f"step {step}")
    assert tb.weights_cache_bytes() == before                     # nothing was kept
    # registered: uploaded once, kept, freed on unregister
    tb.register_weights(A)
    for _ in range(2):
        Cm = np.full((n, m), np.nan, np.float32)
        assert tb.llamafile_sgemm(m, n, kb, A, kb, Bq, kb, Cm, m, 0, 1, wt, vdt, H.F32)
    held = tb.weights_cache_bytes()
    assert held > before
    tb.unregister_weights(A)
    assert tb.weights_cache_bytes() == before


def test_llamafile_sgemm_concurrent_callers(tb, oracle):
    """Two host threads (two ggml thread pools / server slots) calling with ith == 0 at once: the per-device lock serialises them."""
    import threading
    m, K, n = 128, 1024, 2
    kb = K // 32
    jobs = []
    for j in range(2):
        W = H.make_weights(H.Q4_0, m, K, seed=20 + j)
        X = H.make_acts(n, K, seed=30 + j)
        Bq = oracle.from_float(X, H.Q8_0)
        ok, Cr = oracle.sgemm(m, n, kb, W, kb, Bq, kb, m, H.Q4_0, H.Q8_0)
        jobs.append((W, Bq, Cr))
    errs = []

    def worker(j):
        W, Bq, Cr = jobs[j]
        for _ in range(40):
            Cm = np.full((n, m), np.nan, np.float32)
            if not tb.llamafile_sgemm(m, n, kb, W, kb, Bq, kb, Cm, m, 0, 1, H.Q4_0, H.Q8_0, H.F32):
                errs.append("false")
                return
            if H.scaled_err(Cm, Cr) > H.SCALED_TOL or np.isnan(Cm).any():
                errs.append(f"thread {j}: wrong result")
                return
    ts = [threading.Thread(target=worker, args=(j,)) for j in range(2)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errs, errs
