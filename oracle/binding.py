"""TEST INFRASTRUCTURE — ctypes bindings for the CPU oracle (tb_oracle.c) and, when it has been
built, the reference's own compiled hot path (oracle/_ref/libtb200_ref.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module. The product package (llamafile_b200) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
ORACLE_SO = HERE / "_build" / "liboracle.so"
REF_SO = HERE / "_ref" / "libtb200_ref.so"

F32, F16, Q4_0, Q8_0, Q4_K, Q6_K, Q8_K, I32 = 0, 1, 2, 8, 12, 14, 15, 26
Q4_1, Q5_0, Q8_1, Q5_K, IQ4_XS = 3, 6, 9, 13, 23
TYPE_SIZE = {F32: 4, F16: 2, Q4_0: 18, Q8_0: 34, Q4_K: 144, Q6_K: 210, Q8_K: 292, I32: 4, Q4_1: 20, Q5_0: 22, Q8_1: 36, Q5_K: 176, IQ4_XS: 136}
BLCK = {F32: 1, F16: 1, Q4_0: 32, Q8_0: 32, Q4_K: 256, Q6_K: 256, Q8_K: 256, I32: 1, Q4_1: 32, Q5_0: 32, Q8_1: 32, Q5_K: 256, IQ4_XS: 256}
VEC_DOT_TYPE = {Q4_0: Q8_0, Q8_0: Q8_0, Q4_K: Q8_K, Q6_K: Q8_K, F16: F16, Q4_1: Q8_1, Q5_0: Q8_0, Q5_K: Q8_K, IQ4_XS: Q8_K}   # ggml.c type_traits.vec_dot_type


def row_bytes(t: int, k: int) -> int:
    return k // BLCK[t] * TYPE_SIZE[t]


def build_oracle(force: bool = False) -> Path:
    src = HERE / "tb_oracle.c"
    if force or not ORACLE_SO.exists() or ORACLE_SO.stat().st_mtime < max(src.stat().st_mtime, (HERE / "tb_oracle.h").stat().st_mtime):
        ORACLE_SO.parent.mkdir(exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-fopenmp", "-o", str(ORACLE_SO), str(src), "-lm"])
    return ORACLE_SO


def build_ref() -> Path | None:
    """Runs oracle/build_ref.sh (no-op when /root/reference is absent, e.g. on the GPU box)."""
    subprocess.check_call(["bash", str(HERE / "build_ref.sh")])
    return REF_SO if REF_SO.exists() else None


class Tensor(C.Structure):
    """Mirrors oracle_tensor / ref_tensor_desc: the ggml_tensor fields MixMul reads."""
    _fields_ = [("type", C.c_int), ("ne", C.c_int64 * 4), ("nb", C.c_size_t * 4), ("data", C.c_void_p)]

    @staticmethod
    def of(arr: np.ndarray, type_: int, ne, nb=None) -> "Tensor":
        ne = list(ne) + [1] * (4 - len(ne))
        if nb is None:
            nb = [TYPE_SIZE[type_], row_bytes(type_, ne[0])]
            nb.append(nb[1] * ne[1])
            nb.append(nb[2] * ne[2])
        t = Tensor()
        t.type = type_
        t.ne[:] = ne
        t.nb[:] = nb
        t.data = arr.ctypes.data
        t._keep = arr
        return t


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


class _Common:
    """Shared numpy-level API over either library (prefix 'oracle_' or 'ref_')."""

    def sgemm(self, m, n, k, A, lda, B, ldb, ldc, Atype, Btype, Ctype=F32, nth=1):
        """C[ldc*j + i]; returns (serviced, C[n, ldc] float32, NaN-poisoned where unwritten)."""
        Cm = np.full((n, ldc), np.nan, np.float32)
        ok = self._sgemm(m, n, k, A, lda, B, ldb, Cm, ldc, Atype, Btype, Ctype, nth)
        return bool(ok), Cm

    def mul_mat(self, W: np.ndarray, wtype: int, X: np.ndarray, nth=1):
        """ggml_compute_forward_mul_mat (ggml.c:12277-12403) on contiguous tensors:
        W packed [N rows of K], X float32 [T, K] -> float32 [T, N]."""
        X = _f32(X)
        T, K = X.shape
        N = W.size // row_bytes(wtype, K)
        kb = K // BLCK[wtype]
        ok, out = self.sgemm(N, T, kb, W, kb, X, K, N, wtype, F32, nth=nth)     # :12318-12343
        if ok:
            return out
        vdt = VEC_DOT_TYPE[wtype]
        B = self.from_float(X, vdt)                                            # :12345-12373
        ok, out = self.sgemm(N, T, kb, W, kb, B, kb, N, wtype, vdt, nth=nth)    # :12382-12403
        assert ok, "reference would fall back to ggml's vec_dot loop"
        return out

    def from_float(self, X, vdt):
        X = _f32(X)
        if vdt == Q8_0:
            return self.quantize_q8_0(X)
        if vdt == Q8_K:
            return self.quantize_q8_K(X)
        if vdt == F16:
            return self.fp32_to_fp16(X)
        if vdt == Q8_1:
            return self.quantize_q8_1(X)
        raise ValueError(vdt)


class Oracle(_Common):
    def __init__(self):
        self.lib = L = C.CDLL(str(build_oracle()))
        L.oracle_vec_dot.restype = C.c_float
        L.oracle_mixmul_needs.restype = C.c_size_t
        L.oracle_sgemm.argtypes = [C.c_long] * 3 + [C.c_void_p, C.c_long, C.c_void_p, C.c_long, C.c_void_p, C.c_long] + [C.c_int] * 5
        L.oracle_mixmul_iqk.argtypes = [C.c_long, C.c_long, C.c_long, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_long, C.c_long, C.c_void_p]
        L.oracle_mixmul.argtypes = [C.c_void_p] * 5 + [C.c_size_t]
        for f in ("oracle_quantize_row_q8_0", "oracle_quantize_row_q8_0_ref", "oracle_quantize_row_q8_K", "oracle_quantize_row_q4_0_ref"):
            getattr(L, f).argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        L.oracle_fp32_to_fp16_row.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        L.oracle_fp16_to_fp32_row.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        L.oracle_dequantize_row.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64]
        L.oracle_vec_dot.argtypes = [C.c_int, C.c_int64, C.c_void_p, C.c_void_p]
        for f in ("oracle_blockdot_q4_0_q8_0", "oracle_blockdot_q8_0_q8_0", "oracle_blockdot_q4_K_q8_K", "oracle_blockdot_q6_K_q8_K"):
            getattr(L, f).argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]

    name = "oracle"

    def _rows(self, fn, X, out_type):
        X = _f32(X)
        rows, k = X.reshape(-1, X.shape[-1]).shape
        out = np.zeros(rows * row_bytes(out_type, k), np.uint8)
        fn(_p(X), _p(out), rows * k)   # rows are contiguous, so one call over rows*k elements
        return out

    def quantize_q8_0(self, X): return self._rows(self.lib.oracle_quantize_row_q8_0, X, Q8_0)
    def quantize_q8_0_ref(self, X): return self._rows(self.lib.oracle_quantize_row_q8_0_ref, X, Q8_0)
    def quantize_q8_K(self, X): return self._rows(self.lib.oracle_quantize_row_q8_K, X, Q8_K)
    def quantize_q4_0(self, X): return self._rows(self.lib.oracle_quantize_row_q4_0_ref, X, Q4_0)

    def quantize_q8_1(self, X):
        self.lib.oracle_quantize_row_q8_1.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        return self._rows(self.lib.oracle_quantize_row_q8_1, X, Q8_1)

    def fp32_to_fp16(self, X):
        X = _f32(X); out = np.zeros(X.size, np.uint16)
        self.lib.oracle_fp32_to_fp16_row(_p(X), _p(out), X.size)
        return out.view(np.uint8)

    def dequantize(self, type_, data, n):
        out = np.zeros(n, np.float32)
        self.lib.oracle_dequantize_row(type_, _p(data), _p(out), n)
        return out

    def blockdot(self, xtype, x, y, nb):
        fn = {Q4_0: self.lib.oracle_blockdot_q4_0_q8_0, Q8_0: self.lib.oracle_blockdot_q8_0_q8_0, Q4_K: self.lib.oracle_blockdot_q4_K_q8_K,
              Q6_K: self.lib.oracle_blockdot_q6_K_q8_K}[xtype]
        out = np.zeros(nb * (2 if xtype == Q4_K else 1), np.int32)
        fn(_p(x), _p(y), nb, _p(out))
        return out

    def vec_dot(self, xtype, n, x, y):
        return float(self.lib.oracle_vec_dot(xtype, n, _p(x), _p(y)))

    def _sgemm(self, m, n, k, A, lda, B, ldb, Cm, ldc, Atype, Btype, Ctype, nth):
        return self.lib.oracle_sgemm(m, n, k, _p(A), lda, _p(B), ldb, _p(Cm), ldc, 0, 1, Atype, Btype, Ctype)

    def rms_norm_mul(self, x, w, eps):
        x, w = _f32(x), _f32(w); y = np.empty_like(x)
        self.lib.oracle_rms_norm_mul.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_int64, C.c_void_p]
        self.lib.oracle_rms_norm_mul(_p(x), _p(w), eps, x.size, _p(y))
        return y

    def silu_mul(self, gate, up):
        gate, up = _f32(gate), _f32(up); y = np.empty_like(gate)
        self.lib.oracle_silu_mul.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        self.lib.oracle_silu_mul(_p(gate), _p(up), gate.size, _p(y))
        return y

    def moe_combine(self, experts, weights):
        """llama.cpp/src/llama.cpp:8303-8318 restated: rounded products, added in expert order (plain IEEE fp32)."""
        experts, weights = _f32(experts), _f32(weights)
        acc = experts[:, 0, :] * weights[:, 0:1]
        for j in range(1, experts.shape[1]):
            acc = acc + experts[:, j, :] * weights[:, j:j + 1]
        return acc.astype(np.float32)

    def mixmul_needs(self, w: Tensor, x: Tensor, plan: Tensor) -> int:
        return int(self.lib.oracle_mixmul_needs(C.byref(w), C.byref(x), C.byref(plan)))

    def mixmul(self, w: Tensor, x: Tensor, plan: Tensor, out: Tensor, nth=1) -> bool:
        need = self.mixmul_needs(w, x, plan)
        scratch = np.zeros(need + 8192, np.uint8)
        return bool(self.lib.oracle_mixmul(C.byref(w), C.byref(x), C.byref(plan), C.byref(out), _p(scratch), scratch.size))

    def mixmul_iqk(self, Nx, Ny, ne00, ne11, typeA, A, B, Cm, nb1, nb2, row_mapping, nth=1) -> bool:
        return bool(self.lib.oracle_mixmul_iqk(Nx, Ny, ne00, ne11, typeA, _p(A), _p(B), _p(Cm), nb1, nb2, _p(row_mapping)))


class Reference(_Common):
    """The reference's own compiled code. Raises FileNotFoundError when oracle/_ref was not built."""

    def __init__(self):
        if not REF_SO.exists():
            raise FileNotFoundError(str(REF_SO))
        self.lib = L = C.CDLL(str(REF_SO), mode=C.RTLD_LOCAL)
        L.ref_variant_name.restype = C.c_char_p
        L.ref_mixmul_needs.restype = C.c_size_t
        L.ref_sgemm.argtypes = [C.c_long] * 3 + [C.c_void_p, C.c_long, C.c_void_p, C.c_long, C.c_void_p, C.c_long] + [C.c_int] * 4
        L.ref_mixmul.argtypes = [C.c_void_p] * 5 + [C.c_size_t, C.c_int]
        L.ref_mixmul_iqk.argtypes = [C.c_long, C.c_long, C.c_long, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_long, C.c_long, C.c_void_p, C.c_int]
        for f in ("ref_quantize_q8_0", "ref_quantize_q8_K"):
            getattr(L, f).argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_long, C.c_int]
        for f in ("ref_quantize_q8_0_ref", "ref_quantize_q4_0_ref", "ref_quantize_q4_K_ref", "ref_quantize_q6_K_ref"):
            getattr(L, f).argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_long]
        for f in ("ref_fp32_to_fp16_row", "ref_fp16_to_fp32_row", "ref_dequantize_q4_0", "ref_dequantize_q8_0", "ref_dequantize_q4_K", "ref_dequantize_q6_K"):
            getattr(L, f).argtypes = [C.c_void_p, C.c_void_p, C.c_long]
        for f in ("ref_vec_dot_q4_0_q8_0", "ref_vec_dot_q8_0_q8_0", "ref_vec_dot_q4_K_q8_K", "ref_vec_dot_q6_K_q8_K"):
            getattr(L, f).argtypes = [C.c_long, C.c_void_p, C.c_void_p]
            getattr(L, f).restype = C.c_float
        self.variant = L.ref_variant_name().decode()
        self.max_threads = int(L.ref_max_threads())

    name = "reference"

    def _rows(self, fn, X, out_type, *extra):
        X = _f32(X)
        rows, k = X.reshape(-1, X.shape[-1]).shape
        out = np.zeros(rows * row_bytes(out_type, k), np.uint8)
        fn(_p(X), _p(out), rows, k, *extra)
        return out

    def quantize_q8_0(self, X, nth=1): return self._rows(self.lib.ref_quantize_q8_0, X, Q8_0, nth)
    def quantize_q8_0_ref(self, X): return self._rows(self.lib.ref_quantize_q8_0_ref, X, Q8_0)
    def quantize_q8_K(self, X, nth=1): return self._rows(self.lib.ref_quantize_q8_K, X, Q8_K, nth)
    def quantize_q4_0(self, X): return self._rows(self.lib.ref_quantize_q4_0_ref, X, Q4_0)
    def quantize_q4_K(self, X): return self._rows(self.lib.ref_quantize_q4_K_ref, X, Q4_K)
    def quantize_q6_K(self, X): return self._rows(self.lib.ref_quantize_q6_K_ref, X, Q6_K)

    def quantize_type(self, type_, X):
        """quantize_row_<type>_ref for the §8 f3 types (Q4_1, Q5_0, Q5_K, IQ4_XS)."""
        X = _f32(X)
        rows, k = X.reshape(-1, X.shape[-1]).shape
        out = np.zeros(rows * row_bytes(type_, k), np.uint8)
        self.lib.ref_quantize_type_ref.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_long, C.c_long]
        assert self.lib.ref_quantize_type_ref(type_, _p(X), _p(out), rows, k)
        return out

    def quantize_q8_1(self, X, nth=1):
        self.lib.ref_quantize_q8_1.argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_long, C.c_int]
        return self._rows(self.lib.ref_quantize_q8_1, X, Q8_1, nth)

    def fp32_to_fp16(self, X):
        X = _f32(X); out = np.zeros(X.size, np.uint16)
        self.lib.ref_fp32_to_fp16_row(_p(X), _p(out), X.size)
        return out.view(np.uint8)

    def dequantize(self, type_, data, n):
        out = np.zeros(n, np.float32)
        if type_ in (Q4_1, Q5_0, Q5_K, IQ4_XS):
            self.lib.ref_dequantize_type.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_long]
            assert self.lib.ref_dequantize_type(type_, _p(data), _p(out), n)
            return out
        {Q4_0: self.lib.ref_dequantize_q4_0, Q8_0: self.lib.ref_dequantize_q8_0, Q4_K: self.lib.ref_dequantize_q4_K, Q6_K: self.lib.ref_dequantize_q6_K,
         F16: self.lib.ref_fp16_to_fp32_row}[type_](_p(data), _p(out), n)
        return out

    def vec_dot(self, xtype, n, x, y):
        fn = {Q4_0: self.lib.ref_vec_dot_q4_0_q8_0, Q8_0: self.lib.ref_vec_dot_q8_0_q8_0, Q4_K: self.lib.ref_vec_dot_q4_K_q8_K,
              Q6_K: self.lib.ref_vec_dot_q6_K_q8_K}[xtype]
        return float(fn(n, _p(x), _p(y)))

    def _sgemm(self, m, n, k, A, lda, B, ldb, Cm, ldc, Atype, Btype, Ctype, nth):
        return self.lib.ref_sgemm(m, n, k, _p(A), lda, _p(B), ldb, _p(Cm), ldc, nth, Atype, Btype, Ctype)

    def rms_norm_mul(self, x, w, eps):
        x, w = _f32(x), _f32(w); y = np.empty_like(x)
        self.lib.ref_rms_norm_mul.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_long, C.c_void_p]
        self.lib.ref_rms_norm_mul(_p(x), _p(w), eps, x.size, _p(y))
        return y

    def silu_mul(self, gate, up):
        gate, up = _f32(gate), _f32(up); y = np.empty_like(gate)
        self.lib.ref_silu_mul.argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_void_p]
        self.lib.ref_silu_mul(_p(gate), _p(up), gate.size, _p(y))
        return y

    def moe_combine(self, experts, weights):
        """experts [tokens, used, rows], weights [tokens, used] -> [tokens, rows] through ggml_mul + ggml_add."""
        experts, weights = _f32(experts), _f32(weights)
        tokens, used, rows = experts.shape
        y = np.empty((tokens, rows), np.float32)
        self.lib.ref_moe_combine.argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_long, C.c_long, C.c_void_p]
        self.lib.ref_moe_combine(_p(experts), _p(weights), rows, used, tokens, _p(y))
        return y

    def add(self, a, b):
        a, b = _f32(a), _f32(b); y = np.empty_like(a)
        self.lib.ref_add.argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_void_p]
        self.lib.ref_add(_p(a), _p(b), a.size, _p(y))
        return y

    def mixmul_needs(self, w, x, plan) -> int:
        return int(self.lib.ref_mixmul_needs(C.byref(w), C.byref(x), C.byref(plan)))

    def mixmul(self, w, x, plan, out, nth=1) -> bool:
        need = self.mixmul_needs(w, x, plan)
        scratch = np.zeros(need + 8192, np.uint8)
        return bool(self.lib.ref_mixmul(C.byref(w), C.byref(x), C.byref(plan), C.byref(out), _p(scratch), scratch.size, nth))

    def mixmul_iqk(self, Nx, Ny, ne00, ne11, typeA, A, B, Cm, nb1, nb2, row_mapping, nth=1) -> bool:
        return bool(self.lib.ref_mixmul_iqk(Nx, Ny, ne00, ne11, typeA, _p(A), _p(B), _p(Cm), nb1, nb2, _p(row_mapping), nth))


_singletons: dict = {}


def oracle() -> Oracle:
    if "o" not in _singletons:
        _singletons["o"] = Oracle()
    return _singletons["o"]


def reference() -> Reference | None:
    if "r" not in _singletons:
        try:
            _singletons["r"] = Reference()
        except (FileNotFoundError, OSError):
            _singletons["r"] = None
    return _singletons["r"]
