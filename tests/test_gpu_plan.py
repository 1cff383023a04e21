"""GPU (B200): the decode plan (tb200_plan.cu) — a whole sequence of one-token matmul groups as ONE persistent kernel with
per-warp TMA rings, in-kernel Q8_0 quantization and the fused graph neighbours (RMSNorm*weight, silu(gate)*up, residual
add) — against the oracle, the reference-generated fixtures and the one-shot kernels."""
import sys
import zlib
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


def q4_bytes(N, K, seed, exact=False):
    rng = np.random.default_rng(seed)
    b = rng.integers(0, 256, (N * K // 32, 18), dtype=np.uint8)
    d = np.full(len(b), 2.0 ** -6, np.float16) if exact else (rng.uniform(0.5, 1.5, len(b)) * 2e-3).astype(np.float16)
    b[:, :2] = d.view(np.uint8).reshape(-1, 2)
    return b.reshape(-1)


def run_plan(tb, groups_np, repeats=1, ranges=None, wtype=H.Q4_0):
    """groups_np: list of (list of (W bytes, N), K, x) -> outputs per group/matrix after running the plan."""
    keep, groups = [], []
    for mats, K, x in groups_np:
        ws = [tb.Weights(W, wtype, N, K) for W, N in mats]
        outs = [torch.full((N,), float("nan"), device="cuda") for _, N in mats]
        groups.append({"w": ws, "out": outs, "x": dev(x)})
        keep.append((ws, outs))
    plan = tb.Plan(groups)
    for _ in range(repeats):
        for o in (o for _, outs in keep for o in outs):
            o.fill_(float("nan"))
        if ranges:
            for first, count in ranges:
                plan.run(first, count)
        else:
            plan.run()
        tb.synchronize()
        assert plan.status() == 0, "a plan wait timed out"
    res = [[o.cpu().numpy() for o in outs] for _, outs in keep]
    plan.free()
    for ws, _ in keep:
        for w in ws:
            w.free()
    return res


@pytest.mark.parametrize("K", [256, 1024, 4096, 14336])
def test_plan_single_group_vs_oracle(tb, oracle, K):
    """One group, several matrices with odd row counts (row-pair tails), K from one partial ring stage to 3.5 stages per row."""
    shapes = [301, 64, 1]
    mats = [(q4_bytes(n, K, seed=n + K), n) for n in shapes]
    x = H.make_acts(1, K, seed=K)[0]
    res = run_plan(tb, [(mats, K, x)])
    for (W, n), got in zip(mats, res[0]):
        H.assert_close(got[None], oracle.mul_mat(W, H.Q4_0, x[None]), f"K={K} n={n}")


def test_plan_exact_integer_arithmetic(tb, oracle):
    """Inputs on which every fp32 operation is exact (power-of-two scales, activations that quantize to themselves): the
    in-kernel quantizer, the -8 fold and the in-place high nibbles of the plan kernel must reproduce the oracle bit for bit."""
    K = 2048
    rng = np.random.default_rng(5)
    groups, refs = [], []
    for gi, shapes in enumerate(([777, 130], [2049], [64, 64, 64, 66])):
        x = rng.integers(-12, 13, K).astype(np.float32)
        x[::32] = 127.0
        mats = [(q4_bytes(n, K, seed=gi * 10 + i, exact=True), n) for i, n in enumerate(shapes)]
        groups.append((mats, K, x))
        refs.append([oracle.mul_mat(W, H.Q4_0, x[None])[0] for W, _ in mats])
    res = run_plan(tb, groups, repeats=3)
    for got_g, ref_g in zip(res, refs):
        for got, ref in zip(got_g, ref_g):
            assert np.array_equal(got, ref), float(np.abs(got - ref).max())


def test_plan_many_groups_repeated_and_subranges(tb, oracle):
    """40 groups of mixed shapes (more groups than ring stages, warps with no rows in some groups), replayed several times
    (the arrival counters must come back to zero) and as sub-range launches."""
    rng = np.random.default_rng(9)
    groups, refs = [], []
    for gi in range(40):
        K = int(rng.choice([256, 512, 1024, 3072]))
        shapes = [int(n) for n in rng.integers(1, 700, int(rng.integers(1, 5)))]
        x = H.make_acts(1, K, seed=gi)[0] * float(rng.uniform(0.01, 30))
        mats = [(q4_bytes(n, K, seed=gi * 7 + i), n) for i, n in enumerate(shapes)]
        groups.append((mats, K, x))
        refs.append([oracle.mul_mat(W, H.Q4_0, x[None])[0] for W, _ in mats])
    for ranges in (None, [(0, 1), (1, 2), (3, 17), (20, 20)]):
        res = run_plan(tb, groups, repeats=3, ranges=ranges)
        for gi, (got_g, ref_g) in enumerate(zip(res, refs)):
            for got, ref in zip(got_g, ref_g):
                H.assert_close(got[None], ref[None], f"group {gi} ranges={ranges}")


def test_plan_equals_one_shot_quantizer(tb):
    """The in-kernel activation quantizer is the same arithmetic as quantize_q8_0_kernel: on adversarial rows the plan's result
    equals what the one-shot path computes from the separately quantized activations, within accumulation order."""
    K = 256
    W = q4_bytes(96, K, seed=1)
    w = tb.Weights(W, H.Q4_0, 96, K)
    for row in H.adversarial_rows(K)[:13]:
        if not np.isfinite(row).all() or np.abs(row).max() / 127 > 65504:   # fp16 block scale would overflow to inf on either path
            continue
        out_p = torch.full((96,), float("nan"), device="cuda")
        plan = tb.Plan([{"w": [w], "out": [out_p], "x": dev(row)}])
        plan.run()
        out_1 = torch.full((1, 96), float("nan"), device="cuda")
        tb.mul_mat(w, dev(row[None]), 1, out_1)
        tb.synchronize()
        a, b = out_p.cpu().numpy(), out_1.cpu().numpy()[0]
        assert not np.isnan(a).any()
        den = max(float(np.abs(b).max()), 1e-30)
        assert float(np.abs(a - b).max()) / den < 1e-5
        plan.free()
    w.free()


def test_plan_rejects_unsupported(tb):
    K = 1024
    w8 = tb.Weights(H.make_weights(H.Q8_0, 64, K), H.Q8_0, 64, K)
    out = torch.zeros(64, device="cuda")
    x = torch.zeros(K, device="cuda")
    with pytest.raises(tb.TB200Error):
        tb.Plan([{"w": [w8], "out": [out], "x": x}])          # Q8_0: per-matrix API only
    w4 = tb.Weights(q4_bytes(64, 288, 1), H.Q4_0, 64, 288)
    with pytest.raises(tb.TB200Error):
        tb.Plan([{"w": [w4], "out": [out], "x": x}])          # K % 256 != 0
    w8.free(); w4.free()


# ------------------------------------------------------------------ fused neighbours against the reference's own graph ops
def test_plan_fused_layer_vs_reference_golden(tb, oracle):
    """One transformer layer's decode matmuls with RMSNorm*weight -> qkv, wo + residual, RMSNorm*weight -> gate/up,
    silu(gate)*up -> down + residual, as ONE launch, against fixtures produced by ggml's CPU ops + llamafile_sgemm
    (tests/golden/make_golden_fused.py)."""
    sys.path.insert(0, str(ROOT / "tests" / "golden"))
    import make_golden_fused as M
    g = np.load(ROOT / "tests" / "golden" / "fused_golden.npz")
    Wb = M.layer_weights(oracle.quantize_q4_0)
    assert [zlib.crc32(Wb[k].tobytes()) for k in M.SHAPES] == list(g["crc"])
    W = {k: tb.Weights(Wb[k], H.Q4_0, *M.SHAPES[k]) for k in M.SHAPES}
    x, attn, n1, n2 = (dev(g[f"in_{k}"]) for k in ("x", "attn", "norm1", "norm2"))
    o = {k: torch.full((M.SHAPES[k][0],), float("nan"), device="cuda") for k in ("wq", "wk", "wv", "gate", "up")}
    x2 = torch.full((M.HID,), float("nan"), device="cuda")
    x3 = torch.full((M.HID,), float("nan"), device="cuda")
    plan = tb.Plan([
        {"w": [W["wq"], W["wk"], W["wv"]], "out": [o["wq"], o["wk"], o["wv"]], "x": x, "prologue": tb.PLAN_RMSNORM, "norm_weight": n1, "eps": M.EPS},
        {"w": [W["wo"]], "out": [x2], "x": attn, "residual": [x]},
        {"w": [W["gate"], W["up"]], "out": [o["gate"], o["up"]], "x": x2, "prologue": tb.PLAN_RMSNORM, "norm_weight": n2, "eps": M.EPS},
        {"w": [W["down"]], "out": [x3], "x": o["gate"], "x2": o["up"], "prologue": tb.PLAN_SWIGLU, "residual": [x2]},
    ])
    for rep in range(3):
        plan.run()
        tb.synchronize()
        assert plan.status() == 0
        for k in ("wq", "wk", "wv", "gate", "up"):
            H.assert_close(o[k].cpu().numpy()[None], g[f"out_{k}"][None], f"{k} rep {rep}")
        H.assert_close(x2.cpu().numpy()[None], g["out_x2"][None], "x2 = wo.attn + x")
        H.assert_close(x3.cpu().numpy()[None], g["out_x3"][None], "x3 = down.act + x2")
    plan.free()


def test_plan_llama3_8b_layer_shapes_sampled(tb, oracle):
    """BASELINE config 2 shapes in one plan (K=4096: 4096+1024+1024, 4096, 14336+14336; K=14336: 4096; vocabulary 128256):
    sampled rows against the oracle, every output written."""
    rng = np.random.default_rng(3)
    spec = [([4096, 1024, 1024], 4096), ([4096], 4096), ([14336, 14336], 4096), ([4096], 14336), ([128256], 4096)]
    groups = []
    for gi, (shapes, K) in enumerate(spec):
        x = H.make_acts(1, K, seed=gi)[0]
        groups.append(([(q4_bytes(n, K, seed=gi * 5 + i), n) for i, n in enumerate(shapes)], K, x))
    res = run_plan(tb, groups, repeats=2)
    for (mats, K, x), got_g in zip(groups, res):
        rb = K // 32 * 18
        for (W, n), got in zip(mats, got_g):
            assert not np.isnan(got).any()
            for lo in (0, n // 2 - 7, n - 16):
                ref = oracle.mul_mat(W[lo * rb:(lo + 16) * rb], H.Q4_0, x[None])
                H.assert_close(got[lo:lo + 16][None], ref, f"K={K} n={n} rows {lo}")


# ------------------------------------------------------------------ the neighbours for any token count (tb200_fused.cu)
@pytest.mark.parametrize("tokens,k", [(1, 4096), (5, 512), (33, 14336)])
def test_rms_norm_mul_and_silu_mul_device(tb, oracle, tokens, k):
    rng = np.random.default_rng(tokens + k)
    x = (rng.standard_normal((tokens, k)) * 3).astype(np.float32)
    w = (1 + 0.1 * rng.standard_normal(k)).astype(np.float32)
    y = torch.empty(tokens, k, device="cuda")
    tb.rms_norm_mul(dev(x), dev(w), 1e-5, tokens, k, y)
    tb.synchronize()
    ref = np.stack([oracle.rms_norm_mul(r, w, 1e-5) for r in x])
    got = y.cpu().numpy()
    # bit-exact up to the last bit of the double-precision sum of squares (a tree here, sequential in ggml): allow a 1-ulp scale slip
    assert float(np.abs(got - ref).max() / np.abs(ref).max()) < 2e-7
    assert float((got != ref).mean()) < 0.5
    g, u = (rng.standard_normal((tokens, k)) * 2).astype(np.float32), rng.standard_normal((tokens, k)).astype(np.float32)
    tb.silu_mul(dev(g), dev(u), tokens, k, y)
    tb.synchronize()
    ref = np.stack([oracle.silu_mul(a, b) for a, b in zip(g, u)])
    assert float(np.abs(y.cpu().numpy() - ref).max() / np.abs(ref).max()) < 1e-6


@pytest.mark.parametrize("tokens,k", [(1, 4096), (5, 1000), (33, 64)])
def test_plain_neighbour_ops_device(tb, oracle, tokens, k):
    """ggml_rms_norm, ggml_mul (same shape and a broadcast row), ggml_add, ggml_silu one by one (what the backend runs unfused):
    mul and add are single IEEE operations, hence bit-exact; rms_norm as above; silu within expf's few ulp."""
    rng = np.random.default_rng(7 * tokens + k)
    a = (rng.standard_normal((tokens, k)) * 3).astype(np.float32)
    b = rng.standard_normal((tokens, k)).astype(np.float32)
    row = rng.uniform(0.5, 1.5, k).astype(np.float32)
    y = torch.empty(tokens, k, device="cuda")
    da, db, drow = dev(a), dev(b), dev(row)
    tb.mul(da, db, tokens, k, y); tb.synchronize()
    assert np.array_equal(y.cpu().numpy(), a * b)
    tb.mul(da, drow, tokens, k, y, broadcast_b=True); tb.synchronize()
    assert np.array_equal(y.cpu().numpy(), a * row[None])
    tb.add(da, db, tokens, k, y); tb.synchronize()
    assert np.array_equal(y.cpu().numpy(), a + b)
    tb.rms_norm(da, 1e-6, tokens, k, y); tb.synchronize()
    ref = np.stack([oracle.rms_norm_mul(r, np.ones(k, np.float32), 1e-6) for r in a])
    assert float(np.abs(y.cpu().numpy() - ref).max() / np.abs(ref).max()) < 2e-7
    tb.silu(da, tokens, k, y); tb.synchronize()
    ref = np.stack([oracle.silu_mul(r, np.ones(k, np.float32)) for r in a])
    assert float(np.abs(y.cpu().numpy() - ref).max() / np.abs(ref).max()) < 1e-6


@pytest.mark.parametrize("tokens,used,rows", [(1, 2, 4096), (9, 2, 1000), (4, 1, 64), (3, 4, 300)])
def test_moe_combine_bit_exact(tb, oracle, tokens, used, rows):
    rng = np.random.default_rng(rows)
    e = rng.standard_normal((tokens, used, rows)).astype(np.float32)
    w = rng.uniform(0, 1, (tokens, used)).astype(np.float32)
    out = torch.full((tokens, rows), float("nan"), device="cuda")
    tb.moe_combine(dev(e), dev(w), tokens, used, rows, out)
    tb.synchronize()
    assert np.array_equal(out.cpu().numpy(), oracle.moe_combine(e, w))


@pytest.mark.parametrize("wt", [H.Q4_0, H.Q4_K])
def test_fused_acts_feed_prefill_matmuls(tb, oracle, wt):
    """RMSNorm*weight -> from_float -> matmul and silu(gate)*up -> from_float -> matmul at a prefill token count."""
    T, K, N = 40, 1024, 200
    rng = np.random.default_rng(8)
    x = (rng.standard_normal((T, K)) * 2).astype(np.float32)
    nw = (1 + 0.1 * rng.standard_normal(K)).astype(np.float32)
    W = H.make_weights(wt, N, K, seed=5, realistic_q4k=False)
    w = tb.Weights(W, wt, N, K)
    acts = tb.Acts(wt, K, T)
    out = torch.full((T, N), float("nan"), device="cuda")
    tb.acts_quantize_rmsnorm(acts, dev(x), T, dev(nw), 1e-5)
    tb.mul_mat_acts(w, acts, out)
    tb.synchronize()
    ref = oracle.mul_mat(W, wt, np.stack([oracle.rms_norm_mul(r, nw, 1e-5) for r in x]))
    H.assert_close(out.cpu().numpy(), ref, "rmsnorm -> matmul")
    g, u = (rng.standard_normal((T, K)) * 2).astype(np.float32), rng.standard_normal((T, K)).astype(np.float32)
    tb.acts_quantize_swiglu(acts, dev(g), dev(u), T)
    tb.mul_mat_acts(w, acts, out)
    tb.synchronize()
    ref = oracle.mul_mat(W, wt, np.stack([oracle.silu_mul(a, b) for a, b in zip(g, u)]))
    H.assert_close(out.cpu().numpy(), ref, "swiglu -> matmul")
    w.free(); acts.free()


# ------------------------------------------------------------------ Q4_K plans (Q8_K activations quantized in the kernel)
def q4k_bytes(N, K, seed, exact=False):
    if not exact:
        return H.synth_q4_K(N, K, seed)
    rng = np.random.default_rng(seed)
    nb = N * K // 256
    b = rng.integers(0, 256, (nb, 144), dtype=np.uint8)
    b[:, 0:2] = np.full(nb, 2.0 ** -8, np.float16).view(np.uint8).reshape(-1, 2)    # d
    b[:, 2:4] = np.full(nb, 2.0 ** -7, np.float16).view(np.uint8).reshape(-1, 2)    # dmin
    return b.reshape(-1)


@pytest.mark.parametrize("K", [256, 1024, 4096, 8192, 14336, 28672])
def test_plan_q4_K_single_group_vs_oracle(tb, oracle, K):
    """Q4_K weights: one group, odd row counts, K from one super-block to 7 ring stages per row (the 70B down projection)."""
    shapes = [301, 64, 1] if K <= 8192 else [130, 3]
    mats = [(q4k_bytes(n, K, seed=n + K), n) for n in shapes]
    x = H.make_acts(1, K, seed=K)[0]
    res = run_plan(tb, [(mats, K, x)], wtype=H.Q4_K)
    for (W, n), got in zip(mats, res[0]):
        H.assert_close(got[None], oracle.mul_mat(W, H.Q4_K, x[None]), f"q4_K K={K} n={n}")


def test_plan_q4_K_exact_integer_arithmetic(tb, oracle):
    """Power-of-two d / dmin and activations that Q8_K-quantize to themselves (max |x| = 127 with iscale = -1 exactly... the
    row's maximum is -127, so iscale = 1): every fp32 product is exact up to the final sums, which the oracle forms in the same
    order per super-block; the gate is the usual one, tightened to 1e-6 scaled error."""
    K = 2048
    rng = np.random.default_rng(6)
    groups, refs = [], []
    for gi, shapes in enumerate(([777, 130], [2049], [64, 64, 66])):
        x = rng.integers(-12, 13, K).astype(np.float32)
        x[::256] = -127.0
        mats = [(q4k_bytes(n, K, seed=gi * 10 + i, exact=True), n) for i, n in enumerate(shapes)]
        groups.append((mats, K, x))
        refs.append([oracle.mul_mat(W, H.Q4_K, x[None])[0] for W, _ in mats])
    res = run_plan(tb, groups, repeats=3, wtype=H.Q4_K)
    for got_g, ref_g in zip(res, refs):
        for got, ref in zip(got_g, ref_g):
            assert H.scaled_err(got, ref) <= 1e-6, H.scaled_err(got, ref)


def test_plan_q4_K_equals_one_shot_path(tb):
    """Same arithmetic as the one-shot Q4_K decode kernel fed by quantize_q8_K_kernel: equal within accumulation order, on rows
    with the quantizer's corner cases (all-zero super-block, ties for the maximum, tiny and huge magnitudes)."""
    K, N = 1024, 96
    W = q4k_bytes(N, K, seed=3)
    rng = np.random.default_rng(2)
    x = rng.standard_normal(K).astype(np.float32)
    x[0:256] = 0.0
    x[256] = 3.0; x[300] = -3.0
    x[512:768] *= 1e-20
    x[768:1024] *= 1e18
    w = tb.Weights(W, H.Q4_K, N, K)
    xd = dev(x[None])
    ref = torch.full((1, N), float("nan"), device="cuda")
    tb.mul_mat(w, xd, 1, ref)
    out = torch.full((N,), float("nan"), device="cuda")
    plan = tb.Plan([{"w": [w], "out": [out], "x": xd[0]}])
    plan.run()
    tb.synchronize()
    assert plan.status() == 0
    H.assert_close(out.cpu().numpy()[None], ref.cpu().numpy(), "plan vs one-shot q4_K")
    assert H.scaled_err(out.cpu().numpy(), ref.cpu().numpy()[0]) <= 1e-5
    plan.free(); w.free()


def test_plan_q4_K_chain_with_prologues(tb, oracle):
    """A layer-shaped chain of Q4_K groups with the fused neighbours: RMSNorm -> (a, b) ; SwiGLU(a, b) -> c + residual; repeated
    launches. Checked against the oracle's fp32 neighbours + reference quantizer + mul_mat."""
    K, F = 1024, 2048
    rng = np.random.default_rng(11)
    h = (rng.standard_normal(K) * 0.7).astype(np.float32)
    nw = rng.uniform(0.5, 1.5, K).astype(np.float32)
    Wa, Wb, Wc = q4k_bytes(F, K, 21), q4k_bytes(F, K, 22), q4k_bytes(K, F, 23)
    wa, wb, wc = tb.Weights(Wa, H.Q4_K, F, K), tb.Weights(Wb, H.Q4_K, F, K), tb.Weights(Wc, H.Q4_K, K, F)
    hd, nwd = dev(h), dev(nw)
    a = torch.zeros(F, device="cuda"); b = torch.zeros(F, device="cuda"); c = torch.zeros(K, device="cuda")
    plan = tb.Plan([{"w": [wa, wb], "out": [a, b], "x": hd, "prologue": tb.PLAN_RMSNORM, "norm_weight": nwd, "eps": 1e-5},
                    {"w": [wc], "out": [c], "x": a, "x2": b, "prologue": tb.PLAN_SWIGLU, "residual": [hd]}])
    for _ in range(3):
        plan.run()
        tb.synchronize()
        assert plan.status() == 0
    xn = oracle.rms_norm_mul(h[None], nw, 1e-5)
    ra, rb = oracle.mul_mat(Wa, H.Q4_K, xn)[0], oracle.mul_mat(Wb, H.Q4_K, xn)[0]
    H.assert_close(a.cpu().numpy()[None], ra[None], "q4_K rmsnorm group a")
    H.assert_close(b.cpu().numpy()[None], rb[None], "q4_K rmsnorm group b")
    act = oracle.silu_mul(a.cpu().numpy()[None], b.cpu().numpy()[None])      # from the GPU's own a, b: isolates the second group
    rc = oracle.mul_mat(Wc, H.Q4_K, act)[0] + h
    H.assert_close(c.cpu().numpy()[None], rc[None], "q4_K swiglu + residual group")
    plan.free(); wa.free(); wb.free(); wc.free()


def test_plan_independent_groups(tb, oracle):
    """TB200_PLAN_INDEPENDENT: groups declared independent start without the grid barrier; mixed with dependent groups (one reads
    an independent group's output, so the barrier before IT must still hold) and replayed, for both weight types."""
    for wt, mk in ((H.Q4_0, q4_bytes), (H.Q4_K, q4k_bytes)):
        K, N = 1024, 1024
        rng = np.random.default_rng(17)
        xs = [H.make_acts(1, K, seed=40 + i)[0] for i in range(6)]
        Ws = [mk(N, K, seed=50 + i) for i in range(8)]
        ws = [tb.Weights(W, wt, N, K) for W in Ws]
        outs = [torch.full((N,), float("nan"), device="cuda") for _ in range(8)]
        xd = [dev(x) for x in xs]
        ind = tb.PLAN_QUANT | tb.PLAN_INDEPENDENT
        groups = [{"w": [ws[0]], "out": [outs[0]], "x": xd[0]},
                  {"w": [ws[1]], "out": [outs[1]], "x": xd[1], "prologue": ind},
                  {"w": [ws[2]], "out": [outs[2]], "x": xd[2], "prologue": ind},
                  {"w": [ws[3]], "out": [outs[3]], "x": outs[2]},                       # depends on an independent group's output
                  {"w": [ws[4], ws[5]], "out": [outs[4], outs[5]], "x": xd[3], "prologue": ind},
                  {"w": [ws[6]], "out": [outs[6]], "x": outs[3]},                       # depends across an independent group
                  {"w": [ws[7]], "out": [outs[7]], "x": xd[4], "prologue": ind}]
        plan = tb.Plan(groups)
        for _ in range(4):
            plan.run()
        plan.run(1, 3); plan.run(4, 3)
        tb.synchronize()
        assert plan.status() == 0
        got = [o.cpu().numpy() for o in outs]
        ref_in = [xs[0], xs[1], xs[2], got[2], xs[3], xs[3], got[3], xs[4]]
        for i in range(8):
            H.assert_close(got[i][None], oracle.mul_mat(Ws[i], wt, ref_in[i][None]), f"{H.TYPE_NAME[wt]} independent-groups plan, matrix {i}")
        plan.free()
        for w in ws:
            w.free()
