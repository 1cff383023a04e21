/* TEST INFRASTRUCTURE — not part of the product.
 *
 * The matmul's neighbours in the reference graph, computed by the REFERENCE's own ggml CPU ops
 * (pristine llama.cpp/ggml/src/ggml.c compiled by oracle/build_ref.sh): ggml_rms_norm + ggml_mul (llm_build_norm,
 * llama.cpp/src/llama.cpp:8099-8225) and ggml_silu + ggml_mul (llm_build_ffn, LLM_FFN_SILU + LLM_FFN_PAR). Nothing here
 * computes: each function builds a two-node ggml graph and runs ggml_graph_compute_with_ctx on it. Compiled against the
 * pristine ggml.h so the struct layouts agree with the ggml.c objects it is linked with. */
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "ggml.h"
#include "ggml-backend.h"
#include "ggml-backend-impl.h"
#include "tb200_ggml_backend.h"   /* the product's mirror of llamafile's ggml_backend_api table, filled below with the reference's functions */

#define EXPORT __attribute__((visibility("default")))

static struct ggml_context *ctx_for(long n) {
    struct ggml_init_params ip = {(size_t)n * 4 * 8 + (8u << 20), NULL, false};
    return ggml_init(ip);
}

static void compute(struct ggml_context *ctx, struct ggml_tensor *r, float *y, long n) {
    struct ggml_cgraph *gf = ggml_new_graph(ctx);
    ggml_build_forward_expand(gf, r);
    ggml_graph_compute_with_ctx(ctx, gf, 1);
    memcpy(y, r->data, (size_t)n * 4);
}

/* y = rms_norm(x, eps) * w */
EXPORT void ref_rms_norm_mul(const float *x, const float *w, float eps, long n, float *y) {
    struct ggml_context *ctx = ctx_for(n);
    struct ggml_tensor *a = ggml_new_tensor_1d(ctx, GGML_TYPE_F32, n), *b = ggml_new_tensor_1d(ctx, GGML_TYPE_F32, n);
    memcpy(a->data, x, (size_t)n * 4);
    memcpy(b->data, w, (size_t)n * 4);
    compute(ctx, ggml_mul(ctx, ggml_rms_norm(ctx, a, eps), b), y, n);
    ggml_free(ctx);
}

/* y = silu(gate) * up */
EXPORT void ref_silu_mul(const float *gate, const float *up, long n, float *y) {
    struct ggml_context *ctx = ctx_for(n);
    struct ggml_tensor *a = ggml_new_tensor_1d(ctx, GGML_TYPE_F32, n), *b = ggml_new_tensor_1d(ctx, GGML_TYPE_F32, n);
    memcpy(a->data, gate, (size_t)n * 4);
    memcpy(b->data, up, (size_t)n * 4);
    compute(ctx, ggml_mul(ctx, ggml_silu(ctx, a), b), y, n);
    ggml_free(ctx);
}

/* y = a + b (the residual add after wo / ffn_down) */
EXPORT void ref_add(const float *x, const float *r, long n, float *y) {
    struct ggml_context *ctx = ctx_for(n);
    struct ggml_tensor *a = ggml_new_tensor_1d(ctx, GGML_TYPE_F32, n), *b = ggml_new_tensor_1d(ctx, GGML_TYPE_F32, n);
    memcpy(a->data, x, (size_t)n * 4);
    memcpy(b->data, r, (size_t)n * 4);
    compute(ctx, ggml_add(ctx, a, b), y, n);
    ggml_free(ctx);
}

/* ---------------------------------------------------------------------------------------------
 * The host side of llamafile's DSO handshake for tests: llamafile passes `ggml_backend_api()`
 * (llama.cpp.patches/patches/ggml-backend.c.patch) to ggml_cuda_link(); here the same table is filled with the pristine
 * ggml's functions and handed to tb200_backend_link(), so the TB200 backend is driven by the reference's own
 * ggml-backend.c / ggml-alloc.c machinery (ggml_backend_alloc_ctx_tensors, ggml_backend_tensor_set, ggml_backend_graph_compute).
 * It also proves the mirrored struct layouts at run time: ggml's code dereferences the backend's objects directly.
 * --------------------------------------------------------------------------------------------- */
static bool g_log_disable = false;
static long host_write(int fd, const void *p, long n) { return write(fd, p, (size_t)n); }

EXPORT const struct tb200_ggml_backend_api *ref_backend_api(void) {
    static struct tb200_ggml_backend_api api;
    memset(&api, 0, sizeof api);
    api.FLAG_log_disable = &g_log_disable;
    api.exit = exit, api.free = free, api.malloc = malloc, api.getenv = getenv, api.write = host_write;
    api.ggml_backend_register = (void *)ggml_backend_register;
    api.ggml_backend_buffer_init = (void *)ggml_backend_buffer_init;
    api.ggml_backend_cpu_buffer_from_ptr = (void *)ggml_backend_cpu_buffer_from_ptr;
    api.ggml_backend_cpu_buffer_type = (void *)ggml_backend_cpu_buffer_type;
    api.ggml_backend_buft_get_alloc_size = (void *)ggml_backend_buft_get_alloc_size;
    api.ggml_backend_buft_alloc_buffer = (void *)ggml_backend_buft_alloc_buffer;
    api.ggml_backend_is_cpu = (void *)ggml_backend_is_cpu;
    api.ggml_backend_tensor_get = (void *)ggml_backend_tensor_get;
    api.ggml_backend_tensor_set = (void *)ggml_backend_tensor_set;
    api.ggml_is_quantized = (void *)ggml_is_quantized;
    api.ggml_type_size = (void *)ggml_type_size;
    api.ggml_blck_size = (void *)ggml_blck_size;
    api.ggml_is_transposed = (void *)ggml_is_transposed;
    api.ggml_nbytes = (void *)ggml_nbytes;
    api.ggml_get_unary_op = (void *)ggml_get_unary_op;
    api.ggml_nelements = (void *)ggml_nelements;
    api.ggml_nrows = (void *)ggml_nrows;
    api.ggml_is_permuted = (void *)ggml_is_permuted;
    api.ggml_is_contiguous = (void *)ggml_is_contiguous;
    api.ggml_op_name = (void *)ggml_op_name;
    api.ggml_type_name = (void *)ggml_type_name;
    api.ggml_element_size = (void *)ggml_element_size;
    api.ggml_row_size = (void *)ggml_row_size;
    api.ggml_rope_yarn_corr_dims = (void *)ggml_rope_yarn_corr_dims;
    api.ggml_op_desc = (void *)ggml_op_desc;
    api.ggml_backend_buffer_is_host = (void *)ggml_backend_buffer_is_host;
    api.ggml_guid_matches = (void *)ggml_guid_matches;
    api.ggml_is_empty = (void *)ggml_is_empty;
    api.ggml_backend_buffer_get_usage = (void *)ggml_backend_buffer_get_usage;
    api.ggml_are_same_shape = (void *)ggml_are_same_shape;
    api.ggml_is_contiguous_1 = (void *)ggml_is_contiguous_1;
    api.ggml_is_contiguous_2 = (void *)ggml_is_contiguous_2;
    return &api;
}

/* struct layouts and enum values of the reference, for tests/test_abi.py to compare with include/tb200_ggml_backend.h */
EXPORT int ref_backend_abi(long *out, int cap) {
    const long v[] = {
        GGML_OP_NONE, GGML_OP_MUL_MAT, GGML_OP_MUL_MAT_ID, GGML_OP_RESHAPE, GGML_OP_VIEW, GGML_OP_PERMUTE, GGML_OP_TRANSPOSE,
        GGML_OP_ADD, GGML_OP_MUL, GGML_OP_RMS_NORM, GGML_OP_UNARY, GGML_UNARY_OP_SILU, GGML_TENSOR_FLAG_OUTPUT,
        GGML_STATUS_ALLOC_FAILED, GGML_STATUS_FAILED, GGML_STATUS_SUCCESS, GGML_STATUS_ABORTED,
        GGML_BACKEND_BUFFER_USAGE_ANY, GGML_BACKEND_BUFFER_USAGE_WEIGHTS, GGML_BACKEND_BUFFER_USAGE_COMPUTE,
        offsetof(struct ggml_tensor, type), offsetof(struct ggml_tensor, buffer), offsetof(struct ggml_tensor, ne), offsetof(struct ggml_tensor, nb),
        offsetof(struct ggml_tensor, op), offsetof(struct ggml_tensor, op_params), offsetof(struct ggml_tensor, flags), offsetof(struct ggml_tensor, src),
        offsetof(struct ggml_tensor, view_src), offsetof(struct ggml_tensor, view_offs), offsetof(struct ggml_tensor, data),
        offsetof(struct ggml_cgraph, n_nodes), offsetof(struct ggml_cgraph, nodes),
        sizeof(struct ggml_backend_buffer_type_i), sizeof(struct ggml_backend_buffer_type), offsetof(struct ggml_backend_buffer_type, context),
        sizeof(struct ggml_backend_buffer_i), sizeof(struct ggml_backend_buffer), offsetof(struct ggml_backend_buffer, buft),
        offsetof(struct ggml_backend_buffer, context), offsetof(struct ggml_backend_buffer, size), offsetof(struct ggml_backend_buffer, usage),
        sizeof(struct ggml_backend_i), offsetof(struct ggml_backend_i, graph_compute), offsetof(struct ggml_backend_i, supports_op),
        offsetof(struct ggml_backend_i, supports_buft), offsetof(struct ggml_backend_i, event_synchronize),
        sizeof(struct ggml_backend), offsetof(struct ggml_backend, iface), offsetof(struct ggml_backend, context),
    };
    const int n = (int)(sizeof v / sizeof v[0]);
    for (int i = 0; i < n && i < cap; ++i) out[i] = v[i];
    return n;
}

/* MoE tail (llama.cpp/src/llama.cpp:8303-8318): experts [rows, used, tokens] * weights [1, used, tokens], then the used experts'
 * slices added in order. out [rows, tokens]. */
EXPORT void ref_moe_combine(const float *experts, const float *weights, long rows, long used, long tokens, float *out) {
    struct ggml_context *ctx = ctx_for(rows * used * tokens);
    struct ggml_tensor *e = ggml_new_tensor_3d(ctx, GGML_TYPE_F32, rows, used, tokens), *w = ggml_new_tensor_3d(ctx, GGML_TYPE_F32, 1, used, tokens);
    memcpy(e->data, experts, (size_t)rows * used * tokens * 4);
    memcpy(w->data, weights, (size_t)used * tokens * 4);
    struct ggml_tensor *m = ggml_mul(ctx, e, w), *acc = NULL;
    for (long i = 0; i < used; ++i) {
        struct ggml_tensor *cur = ggml_view_2d(ctx, m, rows, tokens, m->nb[2], i * m->nb[1]);
        acc = i == 0 ? cur : ggml_add(ctx, acc, cur);
    }
    if (used == 1) acc = ggml_cont(ctx, acc);
    compute(ctx, acc, out, rows * tokens);
    ggml_free(ctx);
}
