"""The ggml-backend face of the module (include/tb200_ggml_backend.h) from Python, WITHOUT a ggml in the process: ctypes mirrors of
the interface structs and a small stand-alone host that does what a ggml host does with a backend object - allocate a buffer of
the backend's buffer type, place tensors in it, upload with set_tensor, run graph_compute on a graph of MUL_MAT nodes, read
results back - by calling through the object's own interface tables. bench.py's e2e leg uses it; the tests drive the same
object from the reference's real ggml instead (tests/ggml_host.py)."""
from __future__ import annotations

import ctypes as C

from .lib import BLCK, F32, TYPE_SIZE, TB200Error, lib
from .sgemm import GgmlTensor

vp, sz = C.c_void_p, C.c_size_t
OP_MUL_MAT, OP_MUL_MAT_ID = 23, 24
OP_ADD, OP_MUL, OP_RMS_NORM, OP_UNARY, UNARY_SILU = 2, 6, 20, 63, 10   # include/tb200_ggml_backend.h


class BuftIface(C.Structure):
    _fields_ = [("get_name", C.CFUNCTYPE(C.c_char_p, vp)), ("alloc_buffer", C.CFUNCTYPE(vp, vp, sz)), ("get_alignment", C.CFUNCTYPE(sz, vp)),
                ("get_max_size", vp), ("get_alloc_size", vp), ("is_host", C.CFUNCTYPE(C.c_bool, vp))]


class Buft(C.Structure):
    _fields_ = [("iface", BuftIface), ("context", vp)]


class BufferIface(C.Structure):
    _fields_ = [("get_name", C.CFUNCTYPE(C.c_char_p, vp)), ("free_buffer", C.CFUNCTYPE(None, vp)), ("get_base", C.CFUNCTYPE(vp, vp)),
                ("init_tensor", C.CFUNCTYPE(None, vp, vp)), ("set_tensor", C.CFUNCTYPE(None, vp, vp, vp, sz, sz)),
                ("get_tensor", C.CFUNCTYPE(None, vp, vp, vp, sz, sz)), ("cpy_tensor", vp), ("clear", C.CFUNCTYPE(None, vp, C.c_uint8)),
                ("reset", C.CFUNCTYPE(None, vp))]


class Buffer(C.Structure):
    _fields_ = [("iface", BufferIface), ("buft", vp), ("context", vp), ("size", sz), ("usage", C.c_int)]


class BackendIface(C.Structure):
    _fields_ = [("get_name", C.CFUNCTYPE(C.c_char_p, vp)), ("free", C.CFUNCTYPE(None, vp)), ("get_default_buffer_type", C.CFUNCTYPE(vp, vp)),
                ("set_tensor_async", C.CFUNCTYPE(None, vp, vp, vp, sz, sz)), ("get_tensor_async", C.CFUNCTYPE(None, vp, vp, vp, sz, sz)),
                ("cpy_tensor_async", vp), ("synchronize", C.CFUNCTYPE(None, vp)),
                ("graph_plan_create", vp), ("graph_plan_free", vp), ("graph_plan_update", vp), ("graph_plan_compute", vp),
                ("graph_compute", C.CFUNCTYPE(C.c_int, vp, vp)), ("supports_op", C.CFUNCTYPE(C.c_bool, vp, vp)),
                ("supports_buft", C.CFUNCTYPE(C.c_bool, vp, vp)), ("offload_op", vp),
                ("event_new", vp), ("event_free", vp), ("event_record", vp), ("event_wait", vp), ("event_synchronize", vp)]


class Backend(C.Structure):
    _fields_ = [("guid", vp), ("iface", BackendIface), ("context", vp)]


class CGraph(C.Structure):
    _fields_ = [("size", C.c_int), ("n_nodes", C.c_int), ("n_leafs", C.c_int), ("nodes", C.POINTER(vp)), ("grads", vp), ("leafs", vp),
                ("hash_size", sz), ("hash_used", vp), ("hash_keys", vp), ("order", C.c_int)]


class StandaloneHost:
    """A ggml host in miniature: one backend, one buffer, tensors placed by a bump allocator at the buffer type's alignment."""

    def __init__(self, device: int, buffer_bytes: int):
        L = lib()
        if not L.tb200_backend_link(None):
            raise TB200Error("backend_link: no sm_100 device")
        self.backend_p = L.tb200_backend_init(device)
        if not self.backend_p:
            raise TB200Error("backend_init: " + L.tb200_last_error().decode(errors="replace"))
        self.backend = Backend.from_address(self.backend_p)
        self.buft_p = self.backend.iface.get_default_buffer_type(self.backend_p)
        buft = Buft.from_address(self.buft_p)
        self.align = buft.iface.get_alignment(self.buft_p)
        self.buffer_p = buft.iface.alloc_buffer(self.buft_p, buffer_bytes)
        if not self.buffer_p:
            raise TB200Error("alloc_buffer failed")
        self.buffer = Buffer.from_address(self.buffer_p)
        self.base = self.buffer.iface.get_base(self.buffer_p)
        self.used, self.size = 0, buffer_bytes
        self.tensors, self._graphs = [], []

    def tensor(self, type_: int, ne0: int, ne1: int = 1, op: int = 0, src=()):
        row = ne0 // BLCK[type_] * TYPE_SIZE[type_]
        nbytes = row * ne1
        off = (self.used + self.align - 1) // self.align * self.align
        if off + nbytes > self.size:
            raise TB200Error("stand-alone host buffer exhausted")
        self.used = off + nbytes
        t = GgmlTensor()
        t.type, t.buffer_, t.op_, t.data = type_, self.buffer_p, op, self.base + off
        t.ne[:] = [ne0, ne1, 1, 1]
        t.nb[:] = [TYPE_SIZE[type_], row, nbytes, nbytes]
        for i, s in enumerate(src):
            t.src_[i] = C.addressof(s)
        self.tensors.append(t)
        return t

    def mul_mat(self, a, b):
        """ggml_mul_mat(a, b): a [K, N] weights, b [K, T] f32 -> [N, T] f32."""
        return self.tensor(F32, a.ne[1], b.ne[1], OP_MUL_MAT, (a, b))

    def add(self, a, b): return self.tensor(F32, a.ne[0], a.ne[1], OP_ADD, (a, b))
    def mul(self, a, b): return self.tensor(F32, a.ne[0], a.ne[1], OP_MUL, (a, b))

    def rms_norm(self, a, eps: float):
        t = self.tensor(F32, a.ne[0], a.ne[1], OP_RMS_NORM, (a,))
        t.op_params_[0] = C.c_int32.from_buffer_copy(C.c_float(eps)).value
        return t

    def silu(self, a):
        t = self.tensor(F32, a.ne[0], a.ne[1], OP_UNARY, (a,))
        t.op_params_[0] = UNARY_SILU
        return t

    def set(self, t, ptr: int, nbytes: int):
        self.buffer.iface.set_tensor(self.buffer_p, C.addressof(t), ptr, 0, nbytes)

    def graph(self, nodes):
        arr = (vp * len(nodes))(*[C.addressof(n) for n in nodes])
        g = CGraph()
        g.size = g.n_nodes = len(nodes)
        g.nodes = C.cast(arr, C.POINTER(vp))
        self._graphs.append((g, arr, nodes))
        return g

    def set_async(self, t, ptr, nbytes): self.backend.iface.set_tensor_async(self.backend_p, C.addressof(t), ptr, 0, nbytes)
    def get_async(self, t, ptr, nbytes): self.backend.iface.get_tensor_async(self.backend_p, C.addressof(t), ptr, 0, nbytes)
    def compute(self, g) -> int: return self.backend.iface.graph_compute(self.backend_p, C.addressof(g))
    def synchronize(self): self.backend.iface.synchronize(self.backend_p)

    def stats(self):
        a, b = C.c_int(), C.c_int()
        lib().tb200_backend_last_graph_stats(self.backend_p, C.byref(a), C.byref(b))
        return a.value, b.value

    def free(self):
        if getattr(self, "buffer_p", None):
            self.buffer.iface.free_buffer(self.buffer_p)
            lib().tb200_free  # keep the library alive until here
            self.buffer_p = None
        if getattr(self, "backend_p", None):
            self.backend.iface.free(self.backend_p)
            self.backend_p = None
