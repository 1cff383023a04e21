/* TEST INFRASTRUCTURE — not part of the product.
 *
 * The matmul's neighbours in the reference graph, computed by the REFERENCE's own ggml CPU ops
 * (pristine llama.cpp/ggml/src/ggml.c compiled by oracle/build_ref.sh): ggml_rms_norm + ggml_mul (llm_build_norm,
 * llama.cpp/src/llama.cpp:8099-8225) and ggml_silu + ggml_mul (llm_build_ffn, LLM_FFN_SILU + LLM_FFN_PAR). Nothing here
 * computes: each function builds a two-node ggml graph and runs ggml_graph_compute_with_ctx on it. Compiled against the
 * pristine ggml.h so the struct layouts agree with the ggml.c objects it is linked with. */
#include <string.h>

#include "ggml.h"

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
