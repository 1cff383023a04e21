"""Host-side mirror of the reference's interface for this path (llamafile/sgemm.h:23-28, 76-77):
same names, argument order and true/false behaviour, numpy arrays as the host buffers. Each call
goes through the C-ABI entry point of the same name in libtinyblas_b200.so."""
from __future__ import annotations

import ctypes as C

import numpy as np

from .lib import BLCK, F16, F32, I32, TYPE_SIZE, TB200Error, lib


def _p(a):
    return a.ctypes.data if a is not None else 0


def llamafile_sgemm(m, n, k, A, lda, B, ldb, Cm, ldc, ith, nth, Atype, Btype, Ctype) -> bool:
    """bool llamafile_sgemm(m, n, k, A, lda, B, ldb, C, ldc, ith, nth, Atype, Btype, Ctype)."""
    return bool(lib().llamafile_sgemm(m, n, k, _p(A), lda, _p(B), ldb, _p(Cm), ldc, ith, nth, Atype, Btype, Ctype))


class GgmlTensor(C.Structure):
    """struct ggml_tensor prefix as laid out by llama.cpp/ggml/include/ggml.h:583-618 (tb200_ggml_tensor_view)."""
    _fields_ = [("type", C.c_int32), ("backend_", C.c_int32), ("buffer_", C.c_void_p), ("ne", C.c_int64 * 4),
                ("nb", C.c_size_t * 4), ("op_", C.c_int32), ("op_params_", C.c_int32 * 16), ("flags_", C.c_int32),
                ("grad_", C.c_void_p), ("src_", C.c_void_p * 10), ("view_src_", C.c_void_p), ("view_offs_", C.c_size_t),
                ("data", C.c_void_p), ("name_", C.c_char * 128), ("extra_", C.c_void_p)]

    @staticmethod
    def of(arr: np.ndarray, type_: int, ne, nb=None) -> "GgmlTensor":
        ne = list(ne) + [1] * (4 - len(ne))
        if nb is None:
            nb = [TYPE_SIZE[type_], ne[0] // BLCK[type_] * TYPE_SIZE[type_]]
            nb.append(nb[1] * ne[1])
            nb.append(nb[2] * ne[2])
        t = GgmlTensor()
        t.type = type_
        t.ne[:] = ne
        t.nb[:] = nb
        t.data = arr.ctypes.data
        t._keep = arr
        return t


class GgmlComputeParams(C.Structure):
    """struct ggml_compute_params as patched by llamafile (llama.cpp.patches/patches/ggml.h.patch:48-57)."""
    _fields_ = [("ith", C.c_int), ("nth", C.c_int), ("wsize", C.c_size_t), ("wdata", C.c_void_p), ("shared", C.c_void_p)]


def llamafile_mixmul_needs(weights: GgmlTensor, thought: GgmlTensor, plan: GgmlTensor) -> int:
    return int(lib().llamafile_mixmul_needs(C.addressof(weights), C.addressof(thought), C.addressof(plan)))


def llamafile_mixmul(params: GgmlComputeParams, weights: GgmlTensor, thought: GgmlTensor, plan: GgmlTensor, result: GgmlTensor) -> bool:
    return bool(lib().llamafile_mixmul(C.addressof(params), C.addressof(weights), C.addressof(thought), C.addressof(plan), C.addressof(result)))


def llamafile_mixmul_iqk(Nx, Ny, ne00, ne11, typeA, A, B, Cm, nb1, nb2, row_mapping, ith, nth) -> bool:
    return bool(lib().llamafile_mixmul_iqk(Nx, Ny, ne00, ne11, typeA, _p(A), _p(B), _p(Cm), nb1, nb2, _p(row_mapping), ith, nth))


def quantize_row_q8_0(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, np.float32)
    y = np.zeros(x.size // 32 * 34, np.uint8)
    if not lib().tb200_quantize_row_q8_0(_p(x), _p(y), x.size):
        raise TB200Error("quantize_row_q8_0: " + lib().tb200_last_error().decode(errors="replace"))
    return y


def quantize_row_q8_K(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, np.float32)
    y = np.zeros(x.size // 256 * 292, np.uint8)
    if not lib().tb200_quantize_row_q8_K(_p(x), _p(y), x.size):
        raise TB200Error("quantize_row_q8_K: " + lib().tb200_last_error().decode(errors="replace"))
    return y


def fp32_to_fp16_row(x: np.ndarray) -> np.ndarray:
    """ggml_fp32_to_fp16_row on the device (f32_to_f16_kernel); returns the IEEE half bits as uint16."""
    x = np.ascontiguousarray(x, np.float32)
    y = np.zeros(x.size, np.uint16)
    if not lib().tb200_fp32_to_fp16_row(_p(x), _p(y), x.size):
        raise TB200Error("fp32_to_fp16_row: " + lib().tb200_last_error().decode(errors="replace"))
    return y


def weights_cache_clear():
    lib().tb200_weights_cache_clear()


def register_weights(arr: np.ndarray):
    """Declare a host array immutable (the mmap'd GGUF range): llamafile_sgemm keeps resident copies of matrices inside it."""
    if lib().tb200_register_weights(_p(arr), arr.nbytes) != 0:
        raise TB200Error("register_weights: " + lib().tb200_last_error().decode(errors="replace"))


def unregister_weights(arr: np.ndarray):
    if lib().tb200_unregister_weights(_p(arr)) != 0:
        raise TB200Error("unregister_weights: " + lib().tb200_last_error().decode(errors="replace"))


def weights_cache_bytes() -> int:
    return int(lib().tb200_weights_cache_bytes())
