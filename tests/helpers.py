"""Shared test utilities: synthetic inputs, error metrics, adversarial quantizer rows."""
from __future__ import annotations

import numpy as np

from oracle import binding as B

F32, F16, Q4_0, Q8_0, Q4_K, Q6_K, Q8_K, I32 = B.F32, B.F16, B.Q4_0, B.Q8_0, B.Q4_K, B.Q6_K, B.Q8_K, B.I32
Q4_1, Q5_0, Q8_1, Q5_K, IQ4_XS = B.Q4_1, B.Q5_0, B.Q8_1, B.Q5_K, B.IQ4_XS
WEIGHT_TYPES = [Q4_0, Q8_0, Q4_K, Q6_K, F16]
G8_TYPES = [Q4_1, Q5_0, Q5_K, IQ4_XS]   # served through the generic resident form
TYPE_NAME = {F32: "f32", F16: "f16", Q4_0: "q4_0", Q8_0: "q8_0", Q4_K: "q4_K", Q6_K: "q6_K", Q8_K: "q8_K", Q4_1: "q4_1", Q5_0: "q5_0", Q8_1: "q8_1",
             Q5_K: "q5_K", IQ4_XS: "iq4_xs"}

# fp32 output gate (north_star): max|c-r| / max|r| <= 1e-3, plus the reference's own backend gate NMSE <= 5e-4
# (llama.cpp/tests/test-backend-ops.cpp:966-968)
SCALED_TOL = 1e-3
NMSE_TOL = 5e-4


def scaled_err(c, r) -> float:
    c = np.asarray(c, np.float64); r = np.asarray(r, np.float64)
    den = np.abs(r).max()
    return float(np.abs(c - r).max() / den) if den > 0 else float(np.abs(c - r).max())


def nmse(c, r) -> float:
    c = np.asarray(c, np.float64); r = np.asarray(r, np.float64)
    den = (r * r).sum()
    return float(((c - r) ** 2).sum() / den) if den > 0 else float(((c - r) ** 2).sum())


def assert_close(c, r, what=""):
    assert not np.isnan(c).any(), f"{what}: unwritten (NaN-poisoned) outputs"
    se, nm = scaled_err(c, r), nmse(c, r)
    assert se <= SCALED_TOL and nm <= NMSE_TOL, f"{what}: scaled_err={se:.3e} nmse={nm:.3e}"


def synth_q4_K(rows: int, k: int, seed: int) -> np.ndarray:
    """Random but well-formed block_q4_K bytes (any byte pattern is a valid block; scales kept finite and small)."""
    rng = np.random.default_rng(seed)
    nb = rows * k // 256
    blk = np.zeros((nb, 144), np.uint8)
    d = (rng.uniform(0.5, 2.0, nb) * 1e-3).astype(np.float16)
    dmin = (rng.uniform(0.5, 2.0, nb) * 4e-3).astype(np.float16)
    blk[:, 0:2] = d.view(np.uint8).reshape(nb, 2)
    blk[:, 2:4] = dmin.view(np.uint8).reshape(nb, 2)
    blk[:, 4:] = rng.integers(0, 256, (nb, 140), dtype=np.uint8)
    return blk.reshape(-1)


def synth_q6_K(rows: int, k: int, seed: int) -> np.ndarray:
    """Random but well-formed block_q6_K bytes: any ql/qh pattern and any int8 scale is a valid block; d kept small."""
    rng = np.random.default_rng(seed)
    nb = rows * k // 256
    blk = rng.integers(0, 256, (nb, 210), dtype=np.uint8)
    d = (rng.uniform(0.5, 2.0, nb) * 2e-5).astype(np.float16)
    blk[:, 208:210] = d.view(np.uint8).reshape(nb, 2)
    return blk.reshape(-1)


def synth_g8(wtype: int, rows: int, k: int, seed: int) -> np.ndarray:
    """Random but well-formed blocks of the §8 f3 types (any byte pattern is valid; the fp16 scales kept finite and small)."""
    rng = np.random.default_rng(seed)
    ts, blck = B.TYPE_SIZE[wtype], B.BLCK[wtype]
    nb = rows * k // blck
    blk = rng.integers(0, 256, (nb, ts), dtype=np.uint8)
    d = (rng.uniform(0.5, 2.0, nb) * {Q4_1: 2e-3, Q5_0: 1e-3, Q5_K: 4e-5, IQ4_XS: 1e-5}[wtype]).astype(np.float16)
    blk[:, 0:2] = d.view(np.uint8).reshape(nb, 2)
    if wtype in (Q4_1, Q5_K):
        m = (rng.uniform(-1.0, 1.0, nb) * (2e-2 if wtype == Q4_1 else 2e-4)).astype(np.float16)
        if wtype == Q5_K:
            m = np.abs(m)
        blk[:, 2:4] = m.view(np.uint8).reshape(nb, 2)
    return blk.reshape(-1)


def make_weights(wtype: int, rows: int, k: int, seed: int = 1, realistic_q4k: bool = True) -> np.ndarray:
    """fp32 N(0, 0.02^2) quantized with the reference's *_ref quantizers (oracle restatements). Q4_K uses the
    reference library when present (quantize_row_q4_K_ref is weight-prep, outside the hot path), else
    synthetic blocks."""
    rng = np.random.default_rng(seed)
    o = B.oracle()
    if wtype == Q4_K:
        r = B.reference() if realistic_q4k else None
        if r is None:
            return synth_q4_K(rows, k, seed)
        return r.quantize_q4_K((rng.standard_normal((rows, k)) * 0.02).astype(np.float32))
    if wtype == Q6_K:
        r = B.reference() if realistic_q4k else None
        if r is None:
            return synth_q6_K(rows, k, seed)
        return r.quantize_q6_K((rng.standard_normal((rows, k)) * 0.02).astype(np.float32))
    if wtype in G8_TYPES:
        r = B.reference() if realistic_q4k else None
        if r is None:
            return synth_g8(wtype, rows, k, seed)
        return r.quantize_type(wtype, (rng.standard_normal((rows, k)) * 0.02).astype(np.float32))
    w = (rng.standard_normal((rows, k)) * 0.02).astype(np.float32)
    if wtype == Q4_0:
        return o.quantize_q4_0(w)
    if wtype == Q8_0:
        return o.quantize_q8_0_ref(w)
    if wtype == F16:
        return o.fp32_to_fp16(w)
    raise ValueError(wtype)


def make_acts(tokens: int, k: int, seed: int = 2) -> np.ndarray:
    return np.random.default_rng(seed).standard_normal((tokens, k)).astype(np.float32)


def make_plan(tokens: int, experts: int, thinkers: int, seed: int = 3, skew: bool = False) -> np.ndarray:
    rng = np.random.default_rng(seed)
    if skew:
        return np.tile(np.arange(thinkers, dtype=np.int32), (tokens, 1))
    return np.stack([rng.permutation(experts)[:thinkers] for _ in range(tokens)]).astype(np.int32)


def adversarial_rows(k: int = 256) -> np.ndarray:
    """Quantizer edge cases: zero block, single non-zero, +-max ties (first occurrence decides the
    Q8_K sign), exact .5 products, -0.0, denormals, huge magnitudes (fp16 scale overflow), tiny magnitudes."""
    rows = []
    z = np.zeros(k, np.float32); rows.append(z.copy())
    a = z.copy(); a[5] = 3.25; rows.append(a)
    a = z.copy(); a[7] = -1.0; a[9] = 1.0; rows.append(a)            # tie, negative first
    a = z.copy(); a[7] = 1.0; a[9] = -1.0; rows.append(a)            # tie, positive first
    a = z.copy(); a[k - 1] = -2.0; a[0] = 2.0; a[100 % k] = 2.0; rows.append(a)
    a = np.full(k, 127.0, np.float32); a[1::2] = 0.5; a[2::4] = 1.5; a[3::8] = -2.5; rows.append(a)   # x*id hits k+0.5
    a = np.full(k, -0.0, np.float32); a[3] = 1e-3; rows.append(a)
    a = np.full(k, 1e-41, np.float32); rows.append(a)                # denormals only
    a = np.linspace(-1, 1, k).astype(np.float32) * 1e-30; rows.append(a)
    a = np.linspace(-1, 1, k).astype(np.float32) * 65504.0 * 127.0; rows.append(a)   # d = 65504 (fp16 max)
    a = np.linspace(-1, 1, k).astype(np.float32) * 1e10; rows.append(a)             # d overflows fp16 -> inf
    a = (np.arange(k) % 32 - 15.5).astype(np.float32); rows.append(a)
    a = (0.1 + 2 * np.cos(np.arange(k))).astype(np.float32); rows.append(a)         # test-quantize-fns.cpp:28-32
    rng = np.random.default_rng(7)
    rows.append(rng.standard_normal(k).astype(np.float32))
    rows.append((rng.standard_normal(k) * 1e4).astype(np.float32))
    a = np.round(rng.standard_normal(k) * 20).astype(np.float32); rows.append(a)     # many exact ties
    return np.stack(rows)


def lcg_uniform(n: int, seed: int = 1) -> np.ndarray:
    """llamafile/numba.h:3-28 style LCG uniform(-1, 1) (the reference tests' input generator), vectorised."""
    out = np.empty(n, np.float32)
    x = np.uint64(seed)
    a, c = np.uint64(6364136223846793005), np.uint64(1442695040888963407)
    with np.errstate(over="ignore"):
        for i in range(n):
            x = x * a + c
            out[i] = (float((x >> np.uint64(40)) & np.uint64(0xFFFFFF)) / float(1 << 24)) * 2.0 - 1.0
    return out
