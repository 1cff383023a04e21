"""GPU (B200): tb200_mul_mat_noalloc, the entry point for callers that may not allocate (no counterpart in the reference).
Called through the C ABI and compared with the existing entry point tb200_mul_mat —
same kernels, so bit-identical — and with the oracle inside the usual fp32 gate, on two calls in a row that share the library's
fixed scratch array. Collected last on purpose (file name)."""
import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

SCRATCH_BYTES = 320 * 1024   # TB200_NOALLOC_SCRATCH_BYTES (include/tinyblas_b200.h)


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def both(tb, w, X, rows):
    """(existing entry point, no-alloc entry point) on the same inputs, NaN-poisoned outputs."""
    x = dev(X)
    T = X.shape[0]
    ref = torch.full((T, rows), float("nan"), device="cuda", dtype=torch.float32)
    out = torch.full((T, rows), float("nan"), device="cuda", dtype=torch.float32)
    tb.mul_mat(w, x, T, ref)
    tb.mul_mat_noalloc(w, x, T, out)
    tb.synchronize()
    return ref.cpu().numpy(), out.cpu().numpy()


@pytest.mark.parametrize("wt", [H.Q4_0, H.Q8_0, H.Q4_K, H.Q6_K, H.Q5_0, H.F16])
def test_noalloc_matches_mul_mat_two_calls_in_a_row(tb, oracle, wt):
    N, K1, K2 = 520, 4096, 1024
    W1 = H.make_weights(wt, N, K1, seed=11, realistic_q4k=False)
    W2 = H.make_weights(wt, N, K2, seed=12, realistic_q4k=False)
    w1, w2 = tb.Weights(W1, wt, N, K1), tb.Weights(W2, wt, N, K2)
    try:
        # first call fills most of the shared array (8 tokens x 4096), the second one reads a smaller region of what it left
        for w, W, K, T, seed in ((w1, W1, K1, 8, 5), (w2, W2, K2, 3, 6), (w1, W1, K1, 1, 7)):
            X = H.make_acts(T, K, seed=seed)
            ref, out = both(tb, w, X, N)
            assert not np.isnan(out).any(), "unwritten outputs"
            assert np.array_equal(ref, out), f"{H.TYPE_NAME[wt]} K={K} T={T}: differs from tb200_mul_mat"
            H.assert_close(out, oracle.mul_mat(W, wt, X), f"noalloc {H.TYPE_NAME[wt]} K={K} T={T}")
    finally:
        w1.free(), w2.free()


def test_noalloc_allocates_nothing(tb):
    """Free device memory does not move across calls (outputs preallocated; first call made before measuring: it resolves the
    array's address and loads the kernels)."""
    N, K = 1024, 4096
    w = tb.Weights(H.make_weights(H.Q4_0, N, K, seed=3), H.Q4_0, N, K)
    try:
        xs = [dev(H.make_acts(T, K, seed=T)) for T in (1, 4, 8)]
        outs = [torch.empty((x.shape[0], N), device="cuda", dtype=torch.float32) for x in xs]
        for x, o in zip(xs, outs):
            tb.mul_mat_noalloc(w, x, x.shape[0], o)
        tb.synchronize()
        free0, _ = torch.cuda.mem_get_info()
        for _ in range(3):
            for x, o in zip(xs, outs):
                tb.mul_mat_noalloc(w, x, x.shape[0], o)
        tb.synchronize()
        free1, _ = torch.cuda.mem_get_info()
        assert free1 == free0, f"device memory moved by {free0 - free1} bytes"
    finally:
        w.free()


def test_noalloc_rejects_sizes_beyond_the_fixed_array(tb):
    N = 16
    # (a) more tokens than the decode path takes
    K = 256
    w = tb.Weights(H.make_weights(H.Q8_0, N, K, seed=1), H.Q8_0, N, K)
    x = torch.zeros((9, K), device="cuda")
    out = torch.full((9, N), float("nan"), device="cuda")
    with pytest.raises(tb.TB200Error, match="tokens exceed the limit"):
        tb.mul_mat_noalloc(w, x, 9, out)
    w.free()
    # (b) tokens x k beyond the array: 8 x 65536 x 1.25 bytes = 640 KiB > 320 KiB; 4 tokens of the same matrix fit exactly
    K = 65536
    assert 4 * K * 5 // 4 == SCRATCH_BYTES
    w = tb.Weights(H.make_weights(H.Q8_0, N, K, seed=2), H.Q8_0, N, K)
    x = dev(H.make_acts(8, K, seed=4))
    with pytest.raises(tb.TB200Error, match="the fixed array holds"):
        tb.mul_mat_noalloc(w, x, 8, torch.empty((8, N), device="cuda"))
    tb.synchronize()
    assert np.isnan(out.cpu().numpy()).all(), "a rejected call must not launch anything"
    ref = torch.empty((4, N), device="cuda")
    got = torch.full((4, N), float("nan"), device="cuda")
    tb.mul_mat(w, x, 4, ref)
    tb.mul_mat_noalloc(w, x, 4, got)
    tb.synchronize()
    assert np.array_equal(ref.cpu().numpy(), got.cpu().numpy())
    w.free()
