"""ctypes binding of libtinyblas_b200.so — device-level API (include/tinyblas_b200.h part 2).

Device buffers are passed as integer addresses (e.g. torch.Tensor.data_ptr()); torch is only the
allocator/stream provider, never on the compute path.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import os

PKG = Path(__file__).resolve().parent
# TB200_LIB: an experimental build of the same library (tools/build_variant.sh), for A/B measurements only
LIB_PATH = Path(os.environ["TB200_LIB"]).resolve() if os.environ.get("TB200_LIB") else PKG / "libtinyblas_b200.so"

F32, F16, Q4_0, Q8_0, Q4_K, Q6_K, Q8_K, I32 = 0, 1, 2, 8, 12, 14, 15, 26
Q4_1, Q5_0, Q8_1, Q5_K, IQ4_XS = 3, 6, 9, 13, 23
TYPE_SIZE = {F32: 4, F16: 2, Q4_0: 18, Q8_0: 34, Q4_K: 144, Q6_K: 210, Q8_K: 292, I32: 4, Q4_1: 20, Q5_0: 22, Q8_1: 36, Q5_K: 176, IQ4_XS: 136}
BLCK = {F32: 1, F16: 1, Q4_0: 32, Q8_0: 32, Q4_K: 256, Q6_K: 256, Q8_K: 256, I32: 1, Q4_1: 32, Q5_0: 32, Q8_1: 32, Q5_K: 256, IQ4_XS: 256}
OK, ERR_NO_DEVICE, ERR_UNSUPPORTED, ERR_INVALID, ERR_CUDA = range(5)


class TB200Error(RuntimeError):
    pass


_lib = None


def lib() -> C.CDLL:
    """Loads the CUDA module. Raises (never falls back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise TB200Error(f"{LIB_PATH} not built: run `python -m llamafile_b200.build` (there is no CPU fallback)")
    L = C.CDLL(str(LIB_PATH))
    vp, l, i, sz = C.c_void_p, C.c_long, C.c_int, C.c_size_t
    sig = {
        "tb200_link": (C.c_bool, [vp]), "tb200_device_count": (i, []), "tb200_set_device": (i, [i]),
        "tb200_last_error": (C.c_char_p, []), "tb200_sm_count": (i, []), "tb200_set_stream": (i, [vp]),
        "tb200_synchronize": (i, []), "tb200_launch_count": (C.c_uint64, []),
        "tb200_weights_upload": (vp, [vp, i, l, l, sz]), "tb200_weights_from_device": (vp, [vp, i, l, l, sz]),
        "tb200_weights_free": (None, [vp]), "tb200_weights_bytes": (sz, [vp]),
        "tb200_quantize_q8_0_dev": (i, [vp, l, vp, l, l]), "tb200_quantize_q8_K_dev": (i, [vp, l, vp, l, l]),
        "tb200_mul_mat": (i, [vp, vp, l, l, vp, l]), "tb200_mul_mat_noalloc": (i, [vp, vp, l, l, vp, l]),
        "tb200_sgemm_dev": (i, [vp, l, vp, l, i, vp, l]),
        "tb200_mul_mat_multi": (i, [i, vp, vp, l, l, vp, vp]),
        "tb200_acts_create": (vp, [i, l, l]), "tb200_acts_quantize": (i, [vp, vp, l, l]),
        "tb200_mul_mat_acts": (i, [vp, vp, vp, l]), "tb200_mul_mat_acts_multi": (i, [i, vp, vp, vp, vp]), "tb200_acts_free": (None, [vp]),
        "tb200_ipc_get_handle": (i, [vp, vp]), "tb200_ipc_open_handle": (vp, [vp]), "tb200_ipc_close_handle": (i, [vp]),
        "tb200_peers_create": (vp, [i, i, vp, vp]), "tb200_peers_free": (None, [vp]), "tb200_peers_barrier": (i, [vp]),
        "tb200_peers_timed_out": (i, [vp]), "tb200_memset": (i, [vp, i, sz]),
        "tb200_mul_mat_acts_multi_gather": (i, [vp, i, vp, vp, vp, vp, vp]),
        "tb200_blockdot_dev": (i, [vp, vp, i, vp]), "tb200_mixmul_dev": (i, [vp, i, vp, i, vp, i, l, vp]),
        "tb200_malloc": (vp, [sz]), "tb200_free": (None, [vp]), "tb200_memcpy_h2d": (i, [vp, vp, sz]),
        "tb200_memcpy_d2h": (i, [vp, vp, sz]), "tb200_host_alloc": (vp, [sz]), "tb200_host_free": (None, [vp]),
        "llamafile_sgemm": (C.c_bool, [l, l, l, vp, l, vp, l, vp, l, i, i, i, i, i]),
        "llamafile_mixmul": (C.c_bool, [vp, vp, vp, vp, vp]), "llamafile_mixmul_needs": (sz, [vp, vp, vp]),
        "llamafile_mixmul_iqk": (C.c_bool, [l, l, l, i, i, vp, vp, vp, l, l, vp, i, i]),
        "tb200_quantize_row_q8_0": (C.c_bool, [vp, vp, C.c_int64]), "tb200_quantize_row_q8_K": (C.c_bool, [vp, vp, C.c_int64]),
        "tb200_fp32_to_fp16_row": (C.c_bool, [vp, vp, C.c_int64]), "tb200_weights_cache_clear": (None, []),
        "tb200_register_weights": (i, [vp, sz]), "tb200_unregister_weights": (i, [vp]), "tb200_weights_cache_bytes": (sz, []),
        "tb200_rms_norm_mul_dev": (i, [vp, l, vp, C.c_float, l, l, vp, l]),
        "tb200_rms_norm_dev": (i, [vp, l, C.c_float, l, l, vp, l]), "tb200_mul_dev": (i, [vp, l, vp, l, l, l, vp, l]), "tb200_add_dev": (i, [vp, l, vp, l, l, l, vp, l]),
        "tb200_silu_dev": (i, [vp, l, l, l, vp, l]), "tb200_silu_mul_dev": (i, [vp, vp, l, l, l, vp, l]),
        "tb200_moe_combine_dev": (i, [vp, vp, l, i, l, vp, l]), "tb200_acts_quantize_rmsnorm": (i, [vp, vp, l, l, vp, C.c_float, l]),
        "tb200_acts_quantize_swiglu": (i, [vp, vp, vp, l, l, l]),
        "tb200_plan_create": (vp, [i, vp]), "tb200_plan_create_peers": (vp, [vp, sz, i, vp]), "tb200_plan_run": (i, [vp, i, i]), "tb200_plan_status": (i, [vp]), "tb200_plan_reset": (i, [vp]), "tb200_plan_trace": (l, [vp, vp, l]),
        "tb200_plan_free": (None, [vp]),
        "tb200_backend_link": (C.c_bool, [vp]), "tb200_backend_init": (vp, [i]), "tb200_backend_is_tb200": (C.c_bool, [vp]),
        "tb200_backend_buffer_type": (vp, [i]), "tb200_backend_host_buffer_type": (vp, []), "tb200_backend_split_buffer_type": (vp, [vp]),
        "tb200_backend_reg_devices": (i, []), "tb200_backend_get_device_count": (i, []), "tb200_backend_get_device_memory": (None, [i, vp, vp]),
        "tb200_backend_get_device_description": (None, [i, C.c_char_p, sz]), "tb200_backend_get_device_properties": (None, [i, vp]),
        "tb200_backend_register_host_buffer": (C.c_bool, [vp, sz]), "tb200_backend_unregister_host_buffer": (None, [vp]),
        "tb200_backend_last_graph_stats": (None, [vp, vp, vp]),
        "tb200_gguf_open": (vp, [C.c_char_p]), "tb200_gguf_close": (None, [vp]), "tb200_gguf_tensor_count": (l, [vp]),
        "tb200_gguf_tensor_name": (C.c_char_p, [vp, l]), "tb200_gguf_find": (i, [vp, C.c_char_p, vp, vp, vp, vp]),
        "tb200_gguf_upload": (vp, [vp, C.c_char_p, l, l]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


EXPORTS = None   # filled by tests from include/tinyblas_b200.h


def _check(rc: int, what: str):
    if rc != OK:
        raise TB200Error(f"{what}: status {rc}: {lib().tb200_last_error().decode(errors='replace')}")


def link() -> bool:
    """tb200_link handshake: True iff an sm_100 device is usable."""
    return bool(lib().tb200_link(None))


def _require_device():
    if not link():
        raise TB200Error("no sm_100 (B200) device: tinyBLAS-B200 has no CPU fallback")


def device_count() -> int: return int(lib().tb200_device_count())
def sm_count() -> int: return int(lib().tb200_sm_count())
def launch_count() -> int: return int(lib().tb200_launch_count())
def synchronize(): _check(lib().tb200_synchronize(), "synchronize")
def set_stream(cuda_stream: int): _check(lib().tb200_set_stream(cuda_stream), "set_stream")


def _ptr(x) -> int:
    if x is None:
        return 0
    if isinstance(x, int):
        return x
    if hasattr(x, "data_ptr"):
        return x.data_ptr()
    if hasattr(x, "ctypes"):
        return x.ctypes.data
    raise TypeError(type(x))


class Weights:
    """Resident weight matrix (tb200_weights). `data`: GGUF rows as a numpy uint8 array (host) or a
    torch CUDA uint8 tensor (device); `rows` rows of `k` elements."""

    def __init__(self, data, type_: int, rows: int, k: int, row_stride: int = 0):
        _require_device()
        self.type, self.rows, self.k = type_, rows, k
        on_device = hasattr(data, "is_cuda") and data.is_cuda
        fn = lib().tb200_weights_from_device if on_device else lib().tb200_weights_upload
        self.handle = fn(_ptr(data), type_, rows, k, row_stride)
        if not self.handle:
            raise TB200Error("weights: " + lib().tb200_last_error().decode(errors="replace"))
        if on_device:
            synchronize()

    @property
    def nbytes(self) -> int: return int(lib().tb200_weights_bytes(self.handle))

    def free(self):
        if getattr(self, "handle", None):
            lib().tb200_weights_free(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def mul_mat(w: Weights, x, tokens: int, out, ldx: int | None = None, ldo: int | None = None):
    """out[t, :] = W . vec_dot_type(x[t]) — ggml_compute_forward_mul_mat on device buffers (f32 in/out)."""
    _check(lib().tb200_mul_mat(w.handle, _ptr(x), ldx or w.k, tokens, _ptr(out), ldo or w.rows), "mul_mat")


def mul_mat_noalloc(w: Weights, x, tokens: int, out, ldx: int | None = None, ldo: int | None = None):
    """mul_mat for callers that may not allocate: activation scratch in a fixed device array of the library (header: limits)."""
    _check(lib().tb200_mul_mat_noalloc(w.handle, _ptr(x), ldx or w.k, tokens, _ptr(out), ldo or w.rows), "mul_mat_noalloc")


class Acts:
    """Resident quantized activations (tb200_acts): quantize once, multiply by several weight matrices."""

    def __init__(self, weight_type: int, k: int, max_tokens: int):
        _require_device()
        self.k = k
        self.handle = lib().tb200_acts_create(weight_type, k, max_tokens)
        if not self.handle:
            raise TB200Error("acts: " + lib().tb200_last_error().decode(errors="replace"))

    def quantize(self, x, tokens: int, ldx: int | None = None):
        _check(lib().tb200_acts_quantize(self.handle, _ptr(x), ldx or self.k, tokens), "acts_quantize")

    def free(self):
        if getattr(self, "handle", None):
            lib().tb200_acts_free(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def rms_norm_mul(x, norm_weight, eps: float, tokens: int, k: int, y):
    _check(lib().tb200_rms_norm_mul_dev(_ptr(x), k, _ptr(norm_weight), eps, tokens, k, _ptr(y), k), "rms_norm_mul")


def silu_mul(gate, up, tokens: int, k: int, y):
    _check(lib().tb200_silu_mul_dev(_ptr(gate), _ptr(up), k, tokens, k, _ptr(y), k), "silu_mul")


def rms_norm(x, eps: float, tokens: int, k: int, y):
    _check(lib().tb200_rms_norm_dev(_ptr(x), k, eps, tokens, k, _ptr(y), k), "rms_norm")


def mul(a, b, tokens: int, k: int, y, broadcast_b: bool = False):
    _check(lib().tb200_mul_dev(_ptr(a), k, _ptr(b), 0 if broadcast_b else k, tokens, k, _ptr(y), k), "mul")


def add(a, b, tokens: int, k: int, y, broadcast_b: bool = False):
    _check(lib().tb200_add_dev(_ptr(a), k, _ptr(b), 0 if broadcast_b else k, tokens, k, _ptr(y), k), "add")


def silu(x, tokens: int, k: int, y):
    _check(lib().tb200_silu_dev(_ptr(x), k, tokens, k, _ptr(y), k), "silu")


def moe_combine(expert_out, weights, tokens: int, used: int, rows: int, out):
    _check(lib().tb200_moe_combine_dev(_ptr(expert_out), _ptr(weights), tokens, used, rows, _ptr(out), rows), "moe_combine")


def acts_quantize_rmsnorm(a: Acts, x, tokens: int, norm_weight, eps: float):
    _check(lib().tb200_acts_quantize_rmsnorm(a.handle, _ptr(x), a.k, tokens, _ptr(norm_weight), eps, a.k), "acts_quantize_rmsnorm")


def acts_quantize_swiglu(a: Acts, gate, up, tokens: int):
    _check(lib().tb200_acts_quantize_swiglu(a.handle, _ptr(gate), _ptr(up), a.k, tokens, a.k), "acts_quantize_swiglu")


def mul_mat_acts(w: Weights, a: Acts, out, ldo: int | None = None):
    _check(lib().tb200_mul_mat_acts(w.handle, a.handle, _ptr(out), ldo or w.rows), "mul_mat_acts")


def mul_mat_acts_multi(ws, a: Acts, outs):
    n = len(ws)
    W = (C.c_void_p * n)(*[w.handle for w in ws])
    O = (C.c_void_p * n)(*[_ptr(o) for o in outs])
    Ld = (C.c_long * n)(*[w.rows for w in ws])
    _check(lib().tb200_mul_mat_acts_multi(n, W, a.handle, O, Ld), "mul_mat_acts_multi")


def mul_mat_multi(ws, x, tokens: int, outs, ldx: int | None = None):
    """Decode: several weight matrices sharing K and the activations in ONE launch (wq/wk/wv, gate/up)."""
    n = len(ws)
    W = (C.c_void_p * n)(*[w.handle for w in ws])
    O = (C.c_void_p * n)(*[_ptr(o) for o in outs])
    Ld = (C.c_long * n)(*[w.rows for w in ws])
    _check(lib().tb200_mul_mat_multi(n, W, _ptr(x), ldx or ws[0].k, tokens, O, Ld), "mul_mat_multi")


def sgemm_dev(w: Weights, n: int, B, ldb: int, Btype: int, Cout, ldc: int) -> bool:
    rc = lib().tb200_sgemm_dev(w.handle, n, _ptr(B), ldb, Btype, _ptr(Cout), ldc)
    if rc == ERR_UNSUPPORTED:
        return False
    _check(rc, "sgemm_dev")
    return True


def blockdot(w: Weights, B, Btype: int, out):
    _check(lib().tb200_blockdot_dev(w.handle, _ptr(B), Btype, _ptr(out)), "blockdot")


def quantize_q8_0(x, rows: int, k: int, y, ldx: int | None = None):
    _check(lib().tb200_quantize_q8_0_dev(_ptr(x), ldx or k, _ptr(y), rows, k), "quantize_q8_0")


def quantize_q8_K(x, rows: int, k: int, y, ldx: int | None = None):
    _check(lib().tb200_quantize_q8_K_dev(_ptr(x), ldx or k, _ptr(y), rows, k), "quantize_q8_K")


def mixmul(w: Weights, experts: int, thought, tasks: int, plan, thinkers: int, tokens: int, result) -> bool:
    rc = lib().tb200_mixmul_dev(w.handle, experts, _ptr(thought), tasks, _ptr(plan), thinkers, tokens, _ptr(result))
    if rc == ERR_UNSUPPORTED:
        return False
    _check(rc, "mixmul")
    return True


PLAN_QUANT, PLAN_RMSNORM, PLAN_SWIGLU = 0, 1, 2
PLAN_INDEPENDENT = 0x100


class PlanGroupC(C.Structure):
    """tb200_plan_group (include/tinyblas_b200.h)."""
    _fields_ = [("n", C.c_int), ("prologue", C.c_int), ("w", C.c_void_p * 4), ("out", C.c_void_p * 4), ("residual", C.c_void_p * 4),
                ("x", C.c_void_p), ("x2", C.c_void_p), ("norm_weight", C.c_void_p), ("eps", C.c_float)]


class Plan:
    """Decode plan (tb200_plan): a sequence of one-token matmul groups run as ONE persistent kernel. `groups` is a list of dicts:
    {"w": [Weights...], "out": [buffers...], "x": buffer, optional "prologue", "x2", "norm_weight", "eps", "residual": [buffers or None]}.
    peers: a PeerGroup (one process per GPU) - outputs that point into its symmetric buffer are row slices every rank receives."""

    def __init__(self, groups, peers=None):
        _require_device()
        arr = (PlanGroupC * len(groups))()
        self._keep = groups
        for g, d in zip(arr, groups):
            g.n, g.prologue = len(d["w"]), d.get("prologue", PLAN_QUANT)
            for m, w in enumerate(d["w"]):
                g.w[m], g.out[m] = w.handle, _ptr(d["out"][m])
                res = d.get("residual")
                g.residual[m] = _ptr(res[m]) if res and res[m] is not None else None
            g.x, g.x2, g.norm_weight = _ptr(d["x"]), _ptr(d.get("x2")) or None, _ptr(d.get("norm_weight")) or None
            g.eps = d.get("eps", 0.0)
        self.n = len(groups)
        self.handle = (lib().tb200_plan_create_peers(peers.handle, peers.out_bytes, self.n, arr) if peers is not None
                       else lib().tb200_plan_create(self.n, arr))
        if not self.handle:
            raise TB200Error("plan: " + lib().tb200_last_error().decode(errors="replace"))

    def run(self, first: int = 0, count: int | None = None):
        _check(lib().tb200_plan_run(self.handle, first, self.n - first if count is None else count), "plan_run")

    def status(self) -> int: return int(lib().tb200_plan_status(self.handle))

    def trace(self):
        """First call: switch phase stamps on. Later calls: numpy [n_groups, SMs, 8] int64 of SM-clock stamps (debug aid)."""
        import numpy as np
        n = lib().tb200_plan_trace(self.handle, None, 0)
        if n <= 0:
            return None
        out = np.zeros(n, np.int64)
        lib().tb200_plan_trace(self.handle, out.ctypes.data, n)
        return out.reshape(self.n, -1, 24)
    def reset(self): _check(lib().tb200_plan_reset(self.handle), "plan_reset")

    def free(self):
        if getattr(self, "handle", None):
            lib().tb200_plan_free(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class GgufFile:
    """GGUF tensors straight from an mmap (tb200_gguf_*): metadata of every tensor, raw bytes as numpy views, and
    resident Weights for the matrix types this module holds."""

    def __init__(self, path):
        self.handle = lib().tb200_gguf_open(str(path).encode())
        if not self.handle:
            raise TB200Error("gguf: " + lib().tb200_last_error().decode(errors="replace"))

    def names(self):
        L = lib()
        return [L.tb200_gguf_tensor_name(self.handle, i).decode() for i in range(L.tb200_gguf_tensor_count(self.handle))]

    def info(self, name: str):
        """-> (ggml type, ne[4], host address, bytes)"""
        ty, ne, data, nbytes = C.c_int(), (C.c_int64 * 4)(), C.c_void_p(), C.c_size_t()
        _check(lib().tb200_gguf_find(self.handle, name.encode(), C.byref(ty), ne, C.byref(data), C.byref(nbytes)), "gguf_find")
        return ty.value, list(ne), data.value, nbytes.value

    def raw(self, name: str):
        import numpy as np
        ty, ne, addr, nbytes = self.info(name)
        return np.ctypeslib.as_array((C.c_uint8 * nbytes).from_address(addr)) if nbytes else np.zeros(0, np.uint8)

    def weights(self, name: str, row_lo: int = 0, row_hi: int = 0) -> "Weights":
        _require_device()
        ty, ne, _, _ = self.info(name)
        hnd = lib().tb200_gguf_upload(self.handle, name.encode(), row_lo, row_hi)
        if not hnd:
            raise TB200Error("gguf_upload: " + lib().tb200_last_error().decode(errors="replace"))
        w = Weights.__new__(Weights)
        w.type, w.k, w.handle = ty, ne[0], hnd
        w.rows = (row_hi - row_lo) if row_hi > row_lo else ne[1] * ne[2]
        return w

    def close(self):
        if getattr(self, "handle", None):
            lib().tb200_gguf_close(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---- ggml backend (include/tb200_ggml_backend.h)
def backend_link(api_table: int | None = None) -> bool:
    """tb200_backend_link: api_table = address of the host's ggml_backend_api table, or None for the stand-alone host."""
    return bool(lib().tb200_backend_link(api_table))


def backend_init(device: int = 0) -> int:
    h = lib().tb200_backend_init(device)
    if not h:
        raise TB200Error("backend_init: " + lib().tb200_last_error().decode(errors="replace"))
    return h


def backend_last_graph_stats(backend: int):
    a, b = C.c_int(), C.c_int()
    lib().tb200_backend_last_graph_stats(backend, C.byref(a), C.byref(b))
    return a.value, b.value
