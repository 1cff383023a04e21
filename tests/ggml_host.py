"""TEST INFRASTRUCTURE: a ggml HOST for the TB200 backend, made of the reference's own ggml (pristine ggml.c, ggml-alloc.c,
ggml-backend.c compiled into oracle/_ref/libtb200_ref.so). It plays llamafile's part of the DSO handshake
(llamafile/cuda.c:701-753): hands the module the host's `ggml_backend_api` table, asks it for a backend and a buffer type, and
then lets ggml's own machinery - ggml_backend_alloc_ctx_tensors, ggml_backend_tensor_set/get, ggml_backend_graph_compute - drive
the backend object through its interface table. The same graph is run on ggml's CPU backend for comparison, the way the
reference tests its backends (llama.cpp/tests/test-backend-ops.cpp)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from oracle import binding as B

vp, sz, i64 = C.c_void_p, C.c_size_t, C.c_int64


class InitParams(C.Structure):
    _fields_ = [("mem_size", sz), ("mem_buffer", vp), ("no_alloc", C.c_bool)]


def ggml():
    """The reference's ggml as a ctypes library (None where oracle/_ref is not built)."""
    r = B.reference()
    if r is None:
        return None
    L = r.lib
    if getattr(L, "_host_ready", False):
        return L
    sig = {
        "ggml_init": (vp, [InitParams]), "ggml_free": (None, [vp]), "ggml_tensor_overhead": (sz, []), "ggml_graph_overhead": (sz, []),
        "ggml_new_tensor_1d": (vp, [vp, C.c_int, i64]), "ggml_new_tensor_2d": (vp, [vp, C.c_int, i64, i64]), "ggml_new_tensor_3d": (vp, [vp, C.c_int, i64, i64, i64]),
        "ggml_view_2d": (vp, [vp, vp, i64, i64, sz, sz]),
        "ggml_mul_mat": (vp, [vp, vp, vp]), "ggml_mul_mat_id": (vp, [vp, vp, vp, vp]), "ggml_add": (vp, [vp, vp, vp]),
        "ggml_mul": (vp, [vp, vp, vp]), "ggml_rms_norm": (vp, [vp, vp, C.c_float]), "ggml_silu": (vp, [vp, vp]),
        "ggml_new_graph": (vp, [vp]), "ggml_build_forward_expand": (None, [vp, vp]), "ggml_nbytes": (sz, [vp]),
        "ggml_backend_cpu_init": (vp, []), "ggml_backend_cpu_set_n_threads": (None, [vp, C.c_int]), "ggml_backend_free": (None, [vp]),
        "ggml_backend_alloc_ctx_tensors": (vp, [vp, vp]), "ggml_backend_buffer_free": (None, [vp]), "ggml_backend_buffer_set_usage": (None, [vp, C.c_int]),
        "ggml_backend_tensor_set": (None, [vp, vp, sz, sz]), "ggml_backend_tensor_get": (None, [vp, vp, sz, sz]),
        "ggml_backend_graph_compute": (C.c_int, [vp, vp]), "ggml_backend_supports_op": (C.c_bool, [vp, vp]), "ggml_backend_name": (C.c_char_p, [vp]),
        "ggml_backend_buffer_name": (C.c_char_p, [vp]), "ggml_backend_synchronize": (None, [vp]),
        "ref_backend_api": (vp, []), "ref_backend_abi": (C.c_int, [vp, C.c_int]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    L._host_ready = True
    return L


class Graph:
    """One ggml context (no_alloc) + graph, allocated in a backend's default buffer."""

    def __init__(self, backend, n_tensors=64):
        self.G = ggml()
        self.backend = backend
        self.ctx = self.G.ggml_init(InitParams(self.G.ggml_tensor_overhead() * n_tensors + self.G.ggml_graph_overhead() + 4096, None, True))
        self.buffer = None
        self.gf = None

    def tensor(self, type_, *ne):
        return {1: self.G.ggml_new_tensor_1d, 2: self.G.ggml_new_tensor_2d, 3: self.G.ggml_new_tensor_3d}[len(ne)](self.ctx, type_, *ne)

    def mul_mat(self, a, b): return self.G.ggml_mul_mat(self.ctx, a, b)
    def mul_mat_id(self, as_, b, ids): return self.G.ggml_mul_mat_id(self.ctx, as_, b, ids)
    def add(self, a, b): return self.G.ggml_add(self.ctx, a, b)
    def mul(self, a, b): return self.G.ggml_mul(self.ctx, a, b)
    def rms_norm(self, a, eps): return self.G.ggml_rms_norm(self.ctx, a, eps)
    def silu(self, a): return self.G.ggml_silu(self.ctx, a)

    def build(self, outputs, weights_usage=False):
        self.gf = self.G.ggml_new_graph(self.ctx)
        for t in outputs:
            self.G.ggml_build_forward_expand(self.gf, t)
        self.buffer = self.G.ggml_backend_alloc_ctx_tensors(self.ctx, self.backend)
        assert self.buffer, "ggml_backend_alloc_ctx_tensors failed"
        if weights_usage:
            self.G.ggml_backend_buffer_set_usage(self.buffer, 1)

    def set(self, t, arr):
        arr = np.ascontiguousarray(arr)
        assert arr.nbytes == self.G.ggml_nbytes(t), (arr.nbytes, self.G.ggml_nbytes(t))
        self.G.ggml_backend_tensor_set(t, arr.ctypes.data, 0, arr.nbytes)

    def get(self, t, shape, dtype=np.float32):
        out = np.empty(shape, dtype)
        assert out.nbytes == self.G.ggml_nbytes(t)
        self.G.ggml_backend_tensor_get(t, out.ctypes.data, 0, out.nbytes)
        return out

    def compute(self):
        rc = self.G.ggml_backend_graph_compute(self.backend, self.gf)
        self.G.ggml_backend_synchronize(self.backend)
        return rc

    def free(self):
        if self.buffer:
            self.G.ggml_backend_buffer_free(self.buffer)
        self.G.ggml_free(self.ctx)
        self.buffer = self.ctx = None
