"""CPU: the C-ABI library builds, loads and exports every symbol include/tinyblas_b200.h declares; the struct
views match ggml's layout; without a device the module says so loudly instead of computing on the CPU."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

import helpers as H
from oracle import binding as B

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def tbmod():
    from llamafile_b200.build import build
    build()
    import llamafile_b200 as tb
    return tb


def declared_symbols():
    text = (ROOT / "include" / "tinyblas_b200.h").read_text() + (ROOT / "include" / "tb200_ggml_backend.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"TB200_API\s+[\w\s\*]+?\b(\w+)\s*\(", text)))


def test_header_symbols_exported(tbmod):
    names = declared_symbols()
    assert {"llamafile_sgemm", "llamafile_mixmul", "llamafile_mixmul_needs", "llamafile_mixmul_iqk", "tb200_link",
            "tb200_mul_mat", "tb200_mixmul_dev", "tb200_quantize_q8_0_dev", "tb200_quantize_q8_K_dev", "tb200_plan_create",
            "tb200_backend_link", "tb200_backend_init", "tb200_backend_buffer_type", "tb200_register_weights"} <= set(names)
    raw = C.CDLL(str(tbmod.lib()._name))
    missing = [n for n in names if not hasattr(raw, n)]
    assert not missing, missing


def test_struct_views_match_ggml_layout(tbmod):
    T, P = tbmod.GgmlTensor, tbmod.GgmlComputeParams
    mine = [T.type.offset, T.ne.offset, T.nb.offset, T.data.offset, C.sizeof(T), P.ith.offset, P.nth.offset, P.wsize.offset, P.wdata.offset]
    assert mine[:4] == [0, 16, 48, 256]
    r = B.reference()
    if r is not None:   # this container: check against the reference's own (patched) ggml.h
        out = (C.c_size_t * 9)()
        r.lib.ref_struct_layout(out)
        assert list(out) == mine


def test_no_device_is_loud_not_a_cpu_path(tbmod):
    if tbmod.device_count() > 0:
        pytest.skip("a B200 is present")
    assert tbmod.link() is False
    with pytest.raises(tbmod.TB200Error):
        tbmod.Weights(np.zeros(18 * 8, np.uint8), tbmod.Q4_0, 8, 32)
    with pytest.raises(tbmod.TB200Error):
        tbmod.quantize_row_q8_0(np.zeros(32, np.float32))
    # reference contract: false = "not handled, caller keeps its own loop" (llamafile/sgemm.cpp:104-127); C untouched
    W = H.make_weights(H.Q4_0, 8, 256); X = B.oracle().quantize_q8_0(H.make_acts(1, 256))
    out = np.full((1, 8), np.nan, np.float32)
    assert tbmod.llamafile_sgemm(8, 1, 8, W, 8, X, 8, out, 8, 0, 1, H.Q4_0, H.Q8_0, H.F32) is False
    assert np.isnan(out).all()


@pytest.mark.parametrize("rows,cols,experts,thinkers,tasks,tokens", [(14336, 4096, 8, 2, 1, 512), (4096, 14336, 8, 2, 2, 512),
                                                                     (32, 256, 4, 2, 1, 1), (6144, 16384, 8, 2, 2, 217)])
def test_mixmul_needs_matches_reference_arithmetic(tbmod, rows, cols, experts, thinkers, tasks, tokens):
    """llamafile_mixmul_needs is host arithmetic (tinyblas_cpu_mixmul.inc:97-111): same bytes as the oracle (and the
    reference where built)."""
    dummy = np.zeros(8, np.uint8)
    shapes = ([cols, rows, experts], [cols, tasks, tokens], [thinkers, tokens])
    mine = tbmod.llamafile_mixmul_needs(tbmod.GgmlTensor.of(dummy, H.Q4_0, shapes[0]), tbmod.GgmlTensor.of(dummy, H.F32, shapes[1]),
                                        tbmod.GgmlTensor.of(dummy, H.I32, shapes[2]))
    orc = B.oracle().mixmul_needs(B.Tensor.of(dummy, H.Q4_0, shapes[0]), B.Tensor.of(dummy, H.F32, shapes[1]), B.Tensor.of(dummy, H.I32, shapes[2]))
    assert mine == orc and mine > 0
    r = B.reference()
    if r is not None:
        assert mine == r.mixmul_needs(B.Tensor.of(dummy, H.Q4_0, shapes[0]), B.Tensor.of(dummy, H.F32, shapes[1]), B.Tensor.of(dummy, H.I32, shapes[2]))


def test_c_loader_handshake(tbmod):
    """The plain-C dlopen loader (llamafile/cuda.c pattern) resolves every entry point; ready only with a device."""
    L = C.CDLL(str(ROOT / "llamafile_b200" / "libtb200_loader.so"))
    ready = L.tb200_loader_init(str(tbmod.lib()._name).encode())
    assert ready == (1 if tbmod.device_count() > 0 else 0)
    assert L.tb200_loader_ready() == ready


def test_backend_struct_layouts_match_reference(tmp_path):
    """include/tb200_ggml_backend.h mirrors ggml's backend interface structs and enum values; a probe compiled against OUR header must
    print the numbers the reference's own headers give (oracle/ref_ops.c ref_backend_abi, compiled against llama.cpp/ggml)."""
    import ctypes as C
    import subprocess
    from oracle import binding as B
    r = B.reference()
    if r is None:
        pytest.skip("oracle/_ref not built")
    src = tmp_path / "probe.c"
    src.write_text(r"""
#include <stdio.h>
#include <stddef.h>
#include "tb200_ggml_backend.h"
int main(void) {
    const long v[] = {
        TB200_GGML_OP_NONE, TB200_GGML_OP_MUL_MAT, TB200_GGML_OP_MUL_MAT_ID, TB200_GGML_OP_RESHAPE, TB200_GGML_OP_VIEW, TB200_GGML_OP_PERMUTE, TB200_GGML_OP_TRANSPOSE,
        TB200_GGML_OP_ADD, TB200_GGML_OP_MUL, TB200_GGML_OP_RMS_NORM, TB200_GGML_OP_UNARY, TB200_GGML_UNARY_OP_SILU, TB200_GGML_TENSOR_FLAG_OUTPUT,
        TB200_GGML_STATUS_ALLOC_FAILED, TB200_GGML_STATUS_FAILED, TB200_GGML_STATUS_SUCCESS, TB200_GGML_STATUS_ABORTED,
        TB200_GGML_BUFFER_USAGE_ANY, TB200_GGML_BUFFER_USAGE_WEIGHTS, TB200_GGML_BUFFER_USAGE_COMPUTE,
        offsetof(tb200_ggml_tensor_view, type), offsetof(tb200_ggml_tensor_view, buffer_), offsetof(tb200_ggml_tensor_view, ne), offsetof(tb200_ggml_tensor_view, nb),
        offsetof(tb200_ggml_tensor_view, op_), offsetof(tb200_ggml_tensor_view, op_params_), offsetof(tb200_ggml_tensor_view, flags_), offsetof(tb200_ggml_tensor_view, src_),
        offsetof(tb200_ggml_tensor_view, view_src_), offsetof(tb200_ggml_tensor_view, view_offs_), offsetof(tb200_ggml_tensor_view, data),
        offsetof(tb200_ggml_cgraph_view, n_nodes), offsetof(tb200_ggml_cgraph_view, nodes),
        sizeof(struct tb200_ggml_backend_buffer_type_i), sizeof(struct tb200_ggml_backend_buffer_type), offsetof(struct tb200_ggml_backend_buffer_type, context),
        sizeof(struct tb200_ggml_backend_buffer_i), sizeof(struct tb200_ggml_backend_buffer), offsetof(struct tb200_ggml_backend_buffer, buft),
        offsetof(struct tb200_ggml_backend_buffer, context), offsetof(struct tb200_ggml_backend_buffer, size), offsetof(struct tb200_ggml_backend_buffer, usage),
        sizeof(struct tb200_ggml_backend_i), offsetof(struct tb200_ggml_backend_i, graph_compute), offsetof(struct tb200_ggml_backend_i, supports_op),
        offsetof(struct tb200_ggml_backend_i, supports_buft), offsetof(struct tb200_ggml_backend_i, event_synchronize),
        sizeof(struct tb200_ggml_backend), offsetof(struct tb200_ggml_backend, iface), offsetof(struct tb200_ggml_backend, context),
    };
    for (unsigned i = 0; i < sizeof v / sizeof v[0]; ++i) printf("%ld\n", v[i]);
    return 0;
}
""")
    exe = tmp_path / "probe"
    subprocess.check_call(["gcc", "-I", str(ROOT / "include"), str(src), "-o", str(exe)])
    ours = [int(x) for x in subprocess.check_output([str(exe)], text=True).split()]
    buf = (C.c_long * 128)()
    r.lib.ref_backend_abi.argtypes = [C.c_void_p, C.c_int]
    n = r.lib.ref_backend_abi(buf, 128)
    assert ours == list(buf[:n])
