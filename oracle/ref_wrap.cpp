// TEST INFRASTRUCTURE — not part of the product.
//
// Thin ctypes-friendly wrapper around the REFERENCE's own CPU hot path, linked by
// oracle/build_ref.sh into oracle/_ref/libtb200_ref.so. Nothing here computes: every function
// forwards to an unmodified reference symbol. It exists because
//   * llamafile/sgemm.cpp (the CPUID dispatcher) needs Cosmopolitan's X86_HAVE(); we redo its choice
//     (llamafile/sgemm.cpp:32-82) with __builtin_cpu_supports;
//   * the reference entry points are called by N threads with (ith, nth) and llamafile_mixmul needs
//     a working ggml_barrier(); we drive them the way llamafile/sgemm_matmul_test.cpp:32-40 does
//     (an OpenMP team) and back ggml_barrier with `omp barrier`.
#include <omp.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "llama.cpp/ggml.h"            // PATCHED header (scratch tree): ggml_tensor / ggml_compute_params layout
#include "llamafile/sgemm.h"   // reference prototypes (llamafile/sgemm.h:23-77)

#define EXPORT extern "C" __attribute__((visibility("default")))

// symbols the reference TUs import
extern "C" {
__attribute__((visibility("default"))) bool FLAG_precise = false;      // llamafile/flags.cpp:50 default
__attribute__((visibility("default"))) bool FLAG_log_disable = true;
void ggml_barrier(const struct ggml_compute_params *) {
#pragma omp barrier
}
// llamafile's runtime-dispatched ggml-quants (llama.cpp.patches/llamafile-files/ggml-quants.cpp:782-1597
// picks a per-ISA build of the PATCHED ggml-quants.inc; we link the AVX2 build,
// llamafile-files/ggml-quants-amd-avx2.c — the quantize_row_q8_0 body is the same for every x86 ISA
// level >= AVX, ggml-quants.c:940-990). The patched build matters: llamafile reorders block_q8_K to
// {d, bsums, qs} (llama.cpp.patches/patches/ggml-common.h.patch:25-35).
#define Q(name) name##_amd_avx2
void Q(quantize_row_q8_0)(const float *x, void *y, int64_t k);
void Q(quantize_row_q8_0_ref)(const float *x, void *y, int64_t k);
void Q(quantize_row_q8_K)(const float *x, void *y, int64_t k);
void Q(quantize_row_q4_0_ref)(const float *x, void *y, int64_t k);
void Q(quantize_row_q4_K_ref)(const float *x, void *y, int64_t k);
void Q(quantize_row_q6_K_ref)(const float *x, void *y, int64_t k);
void Q(quantize_row_q4_1_ref)(const float *x, void *y, int64_t k);
void Q(quantize_row_q5_0_ref)(const float *x, void *y, int64_t k);
void Q(quantize_row_q5_K_ref)(const float *x, void *y, int64_t k);
void Q(quantize_row_iq4_xs_ref)(const float *x, void *y, int64_t k);
void Q(quantize_row_q8_1)(const float *x, void *y, int64_t k);
void Q(dequantize_row_q4_1)(const void *x, float *y, int64_t k);
void Q(dequantize_row_q5_0)(const void *x, float *y, int64_t k);
void Q(dequantize_row_q5_K)(const void *x, float *y, int64_t k);
void Q(dequantize_row_iq4_xs)(const void *x, float *y, int64_t k);
void Q(dequantize_row_q4_0)(const void *x, float *y, int64_t k);
void Q(dequantize_row_q8_0)(const void *x, float *y, int64_t k);
void Q(dequantize_row_q4_K)(const void *x, float *y, int64_t k);
void Q(dequantize_row_q6_K)(const void *x, float *y, int64_t k);
void Q(ggml_vec_dot_q4_0_q8_0)(int n, float *s, size_t bs, const void *vx, size_t bx, const void *vy, size_t by, int nrc);
void Q(ggml_vec_dot_q8_0_q8_0)(int n, float *s, size_t bs, const void *vx, size_t bx, const void *vy, size_t by, int nrc);
void Q(ggml_vec_dot_q4_K_q8_K)(int n, float *s, size_t bs, const void *vx, size_t bx, const void *vy, size_t by, int nrc);
void Q(ggml_vec_dot_q6_K_q8_K)(int n, float *s, size_t bs, const void *vx, size_t bx, const void *vy, size_t by, int nrc);
// what llamafile's dispatcher resolves `quantize_row_q8_0` to (imported by tinyblas_cpu_mixmul.inc:343)
void quantize_row_q8_0(const float *x, void *y, int64_t k) { Q(quantize_row_q8_0)(x, y, k); }
}

namespace {

typedef bool sgemm_fn(long, long, long, const void *, long, const void *, long, void *, long, int, int, int, int, int);
typedef bool mixmul_fn(const struct ggml_compute_params *, const struct ggml_tensor *, const struct ggml_tensor *,
                       const struct ggml_tensor *, struct ggml_tensor *);
typedef bool moe_fn(long, long, long, int, int, const void *, const void *, float *, long, long, const void *, int, int);

struct Funcs {
    int variant;
    const char *name;
    sgemm_fn *sgemm;
    mixmul_fn *mixmul;
    moe_fn *iqk_mixmul;
    Funcs() {
        __builtin_cpu_init();
        const char *force = getenv("TB200_REF_VARIANT");   // diagnosis only: "avx2" / "avx512f" force a lower ISA variant than the CPU's
        const bool no512 = force && !strcmp(force, "avx2"), nozen4 = no512 || (force && !strcmp(force, "avx512f"));
        // llamafile/sgemm.cpp:32-60
        if (!nozen4 && __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512vl") &&
            __builtin_cpu_supports("avx512bw") && __builtin_cpu_supports("avx512dq") &&
            __builtin_cpu_supports("avx512vnni") && __builtin_cpu_supports("avx512bf16")) {
            variant = 0, name = "amd_zen4";
            sgemm = llamafile_sgemm_amd_zen4, mixmul = llamafile_mixmul_amd_zen4, iqk_mixmul = iqk_mul_mat_moe_zen4;
        } else if (!no512 && __builtin_cpu_supports("avx512f")) {
            variant = 1, name = "amd_avx512f";
            sgemm = llamafile_sgemm_amd_avx512f, mixmul = llamafile_mixmul_amd_avx512f, iqk_mixmul = iqk_mul_mat_moe;
        } else {
            variant = 2, name = "amd_avx2";
            sgemm = llamafile_sgemm_amd_avx2, mixmul = llamafile_mixmul_amd_avx2, iqk_mixmul = iqk_mul_mat_moe;
        }
        struct ggml_init_params ip = {1 << 20, nullptr, false};
        struct ggml_context *ctx = ggml_init(ip);   // fills ggml_table_f32_f16 (GGML_FP16_TO_FP32 lookups)
        if (ctx) ggml_free(ctx);
    }
};
const Funcs &funcs() { static Funcs f; return f; }

int clamp_threads(int nth) { return nth < 1 ? 1 : nth; }

}  // namespace

EXPORT int ref_variant(void) { return funcs().variant; }
EXPORT const char *ref_variant_name(void) { return funcs().name; }
EXPORT int ref_max_threads(void) { return omp_get_max_threads(); }

// llamafile_sgemm as its callers run it: every thread of a team calls with its own ith.
EXPORT int ref_sgemm(long m, long n, long k, const void *A, long lda, const void *B, long ldb, void *C, long ldc,
                     int nth, int Atype, int Btype, int Ctype) {
    const Funcs &f = funcs();
    int ok = 1;
#pragma omp parallel num_threads(clamp_threads(nth))
    {
        if (!f.sgemm(m, n, k, A, lda, B, ldb, C, ldc, omp_get_thread_num(), omp_get_num_threads(), Atype, Btype, Ctype)) {
#pragma omp atomic write
            ok = 0;
        }
    }
    return ok;
}

struct ref_tensor_desc {
    int type;
    int64_t ne[4];
    size_t nb[4];
    void *data;
};

static void fill(struct ggml_tensor *t, const ref_tensor_desc *d) {
    memset(t, 0, sizeof(*t));
    t->type = (enum ggml_type)d->type;
    for (int i = 0; i < 4; ++i) t->ne[i] = d->ne[i], t->nb[i] = d->nb[i];
    t->data = d->data;
}

EXPORT size_t ref_mixmul_needs(const ref_tensor_desc *w, const ref_tensor_desc *x, const ref_tensor_desc *plan) {
    funcs();
    struct ggml_tensor tw, tx, tp;
    fill(&tw, w), fill(&tx, x), fill(&tp, plan);
    return llamafile_mixmul_needs(&tw, &tx, &tp);
}

EXPORT int ref_mixmul(const ref_tensor_desc *w, const ref_tensor_desc *x, const ref_tensor_desc *plan,
                      const ref_tensor_desc *out, void *wdata, size_t wsize, int nth) {
    const Funcs &f = funcs();
    struct ggml_tensor tw, tx, tp, to;
    fill(&tw, w), fill(&tx, x), fill(&tp, plan), fill(&to, out);
    int ok = 1;
#pragma omp parallel num_threads(clamp_threads(nth))
    {
        struct ggml_compute_params params;
        memset(&params, 0, sizeof(params));
        params.ith = omp_get_thread_num();
        params.nth = omp_get_num_threads();
        params.wsize = wsize;
        params.wdata = wdata;
        if (!f.mixmul(&params, &tw, &tx, &tp, &to)) {
#pragma omp atomic write
            ok = 0;
        }
    }
    return ok;
}

// llamafile_mixmul_iqk (llamafile/sgemm.cpp:139-145) → iqk_mul_mat_moe of the selected ISA
EXPORT int ref_mixmul_iqk(long Nx, long Ny, long ne00, int ne11, int typeA, const void *A, const void *B, float *C,
                          long nb1, long nb2, const void *row_mapping, int nth) {
    const Funcs &f = funcs();
    int ok = 1;
#pragma omp parallel num_threads(clamp_threads(nth))
    {
        if (!f.iqk_mixmul(Nx, Ny, ne00, ne11, typeA, A, B, C, nb1, nb2, row_mapping, omp_get_thread_num(),
                          omp_get_num_threads())) {
#pragma omp atomic write
            ok = 0;
        }
    }
    return ok;
}

// quantizers / converters: rows × k, contiguous
EXPORT void ref_quantize_q8_0(const float *x, void *y, long rows, long k, int nth) {
    funcs();
#pragma omp parallel for num_threads(clamp_threads(nth))
    for (long r = 0; r < rows; ++r) Q(quantize_row_q8_0)(x + r * k, (char *)y + r * (k / 32) * 34, k);
}
EXPORT void ref_quantize_q8_0_ref(const float *x, void *y, long rows, long k) {
    funcs();
    for (long r = 0; r < rows; ++r) Q(quantize_row_q8_0_ref)(x + r * k, (char *)y + r * (k / 32) * 34, k);
}
EXPORT void ref_quantize_q8_K(const float *x, void *y, long rows, long k, int nth) {
    funcs();
#pragma omp parallel for num_threads(clamp_threads(nth))
    for (long r = 0; r < rows; ++r) Q(quantize_row_q8_K)(x + r * k, (char *)y + r * (k / 256) * 292, k);
}
EXPORT void ref_quantize_q4_0_ref(const float *x, void *y, long rows, long k) {
    funcs();
#pragma omp parallel for
    for (long r = 0; r < rows; ++r) Q(quantize_row_q4_0_ref)(x + r * k, (char *)y + r * (k / 32) * 18, k);
}
EXPORT void ref_quantize_q4_K_ref(const float *x, void *y, long rows, long k) {
    funcs();
#pragma omp parallel for
    for (long r = 0; r < rows; ++r) Q(quantize_row_q4_K_ref)(x + r * k, (char *)y + r * (k / 256) * 144, k);
}
EXPORT void ref_quantize_q6_K_ref(const float *x, void *y, long rows, long k) {
    funcs();
#pragma omp parallel for
    for (long r = 0; r < rows; ++r) Q(quantize_row_q6_K_ref)(x + r * k, (char *)y + r * (k / 256) * 210, k);
}
// SURVEY.md §8 f3 types: `type` is the ggml type code; row bytes from the block sizes
EXPORT int ref_quantize_type_ref(int type, const float *x, void *y, long rows, long k) {
    funcs();
    long rb; void (*fn)(const float *, void *, int64_t);
    switch (type) {
    case 3: rb = k / 32 * 20, fn = Q(quantize_row_q4_1_ref); break;
    case 6: rb = k / 32 * 22, fn = Q(quantize_row_q5_0_ref); break;
    case 13: rb = k / 256 * 176, fn = Q(quantize_row_q5_K_ref); break;
    case 23: rb = k / 256 * 136, fn = Q(quantize_row_iq4_xs_ref); break;
    default: return 0;
    }
#pragma omp parallel for
    for (long r = 0; r < rows; ++r) fn(x + r * k, (char *)y + r * rb, k);
    return 1;
}
EXPORT void ref_quantize_q8_1(const float *x, void *y, long rows, long k, int nth) {
    funcs();
#pragma omp parallel for num_threads(clamp_threads(nth))
    for (long r = 0; r < rows; ++r) Q(quantize_row_q8_1)(x + r * k, (char *)y + r * (k / 32) * 36, k);
}
EXPORT int ref_dequantize_type(int type, const void *x, float *y, long n) {
    funcs();
    switch (type) {
    case 3: Q(dequantize_row_q4_1)(x, y, n); return 1;
    case 6: Q(dequantize_row_q5_0)(x, y, n); return 1;
    case 13: Q(dequantize_row_q5_K)(x, y, n); return 1;
    case 23: Q(dequantize_row_iq4_xs)(x, y, n); return 1;
    default: return 0;
    }
}
EXPORT void ref_fp32_to_fp16_row(const float *x, uint16_t *y, long n) { funcs(); ggml_fp32_to_fp16_row(x, y, n); }
EXPORT void ref_fp16_to_fp32_row(const uint16_t *x, float *y, long n) { funcs(); ggml_fp16_to_fp32_row(x, y, n); }
EXPORT void ref_dequantize_q4_0(const void *x, float *y, long n) { funcs(); Q(dequantize_row_q4_0)(x, y, n); }
EXPORT void ref_dequantize_q8_0(const void *x, float *y, long n) { funcs(); Q(dequantize_row_q8_0)(x, y, n); }
EXPORT void ref_dequantize_q4_K(const void *x, float *y, long n) { funcs(); Q(dequantize_row_q4_K)(x, y, n); }
EXPORT void ref_dequantize_q6_K(const void *x, float *y, long n) { funcs(); Q(dequantize_row_q6_K)(x, y, n); }

// the semantic definitions north_star cites (ggml-quants.c:3729, :5227, :6832)
EXPORT float ref_vec_dot_q4_0_q8_0(long n, const void *x, const void *y) {
    funcs(); float s = 0; Q(ggml_vec_dot_q4_0_q8_0)((int)n, &s, 0, x, 0, y, 0, 1); return s;
}
EXPORT float ref_vec_dot_q8_0_q8_0(long n, const void *x, const void *y) {
    funcs(); float s = 0; Q(ggml_vec_dot_q8_0_q8_0)((int)n, &s, 0, x, 0, y, 0, 1); return s;
}
EXPORT float ref_vec_dot_q4_K_q8_K(long n, const void *x, const void *y) {
    funcs(); float s = 0; Q(ggml_vec_dot_q4_K_q8_K)((int)n, &s, 0, x, 0, y, 0, 1); return s;
}

EXPORT float ref_vec_dot_q6_K_q8_K(long n, const void *x, const void *y) {
    funcs(); float s = 0; Q(ggml_vec_dot_q6_K_q8_K)((int)n, &s, 0, x, 0, y, 0, 1); return s;
}

// layout of the two ggml structs llamafile_mixmul reads, as the reference's patched ggml.h defines them
EXPORT void ref_struct_layout(size_t *out) {
    out[0] = offsetof(struct ggml_tensor, type);
    out[1] = offsetof(struct ggml_tensor, ne);
    out[2] = offsetof(struct ggml_tensor, nb);
    out[3] = offsetof(struct ggml_tensor, data);
    out[4] = sizeof(struct ggml_tensor);
    out[5] = offsetof(struct ggml_compute_params, ith);
    out[6] = offsetof(struct ggml_compute_params, nth);
    out[7] = offsetof(struct ggml_compute_params, wsize);
    out[8] = offsetof(struct ggml_compute_params, wdata);
}
