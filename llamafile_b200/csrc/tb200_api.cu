// C-ABI entry points of libtinyblas_b200.so (include/tinyblas_b200.h): module handshake, resident
// weights, device-level matmuls and the reference-compatible host-pointer shims. Host-side C++ only;
// all arithmetic is in the .cu kernels. No CPU fallback: without an sm_100 device every compute entry
// point fails.
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <map>
#include <mutex>
#include <tuple>

#include "tb200_common.cuh"

int tb200_launch_gemm(tb200_ctx *c, const tb200_weights *w, long row0, long nrows, const int8_t *aq, const int2 *ads,
                      const void *af, long a_stride, long tokens, float *out, long ldo, const tb200_peers *peers = nullptr);   // tb200_gemm.cu
bool tb200_gemm_available();
void tb200_gemm_release(tb200_weights *);

namespace {

constexpr int kMaxDevices = 16;
std::mutex g_mu;
tb200_ctx g_ctx[kMaxDevices];
bool g_ctx_init[kMaxDevices];
int g_ndev = -1;
tb200_host_api g_host = {nullptr, nullptr, nullptr, nullptr};
thread_local char g_err[512];
std::atomic<uint64_t> g_launches{0};

void log_line(const char *s) {
    if (g_host.write) g_host.write(2, s, strlen(s));
    else fputs(s, stderr);
}

int probe_devices() {
    if (g_ndev >= 0) return g_ndev;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); n = 0; }
    int usable = 0;
    for (int d = 0; d < n && d < kMaxDevices; ++d) {
        cudaDeviceProp p;
        if (cudaGetDeviceProperties(&p, d) == cudaSuccess && p.major == 10) usable = d + 1;   // sm_100 family only
    }
    g_ndev = usable;
    return g_ndev;
}

size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace

void tb200_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
bool tb200_pdl_enabled() {
    static const int on = [] { const char *v = getenv("TB200_PDL"); return (v && *v == '0') ? 0 : 1; }();
    return on != 0;
}
void tb200_count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

tb200_ctx *tb200_get_ctx() {
    std::lock_guard<std::mutex> lk(g_mu);
    if (probe_devices() <= 0) { tb200_set_error("no sm_100 CUDA device"); return nullptr; }
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev >= g_ndev) { tb200_set_error("cudaGetDevice failed"); return nullptr; }
    tb200_ctx *c = &g_ctx[dev];
    if (!g_ctx_init[dev]) {
        memset(c, 0, sizeof(*c));
        c->device = dev;
        cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, dev);
        c->stream = 0;   // legacy default stream until tb200_set_stream(): orders with any other default-stream user
        g_ctx_init[dev] = true;
    }
    return c;
}

void *tb200_scratch(tb200_ctx *c, size_t bytes) {
    if (bytes <= c->scratch_bytes) return c->scratch;
    cudaStreamSynchronize(c->stream);
    if (c->scratch) cudaFree(c->scratch);
    c->scratch = nullptr, c->scratch_bytes = 0;
    const size_t want = round_up(bytes + bytes / 2, 1 << 20);
    if (cudaMalloc(&c->scratch, want) != cudaSuccess) { tb200_set_error("scratch: cudaMalloc(%zu) failed", want); cudaGetLastError(); return nullptr; }
    c->scratch_bytes = want;
    return c->scratch;
}

void *tb200_scratch2(tb200_ctx *c, size_t bytes) {
    if (bytes <= c->scratch2_bytes) return c->scratch2;
    cudaStreamSynchronize(c->stream);
    if (c->scratch2) cudaFree(c->scratch2);
    c->scratch2 = nullptr, c->scratch2_bytes = 0;
    const size_t want = round_up(bytes + bytes / 2, 1 << 20);
    if (cudaMalloc(&c->scratch2, want) != cudaSuccess) { tb200_set_error("scratch2: cudaMalloc(%zu) failed", want); cudaGetLastError(); return nullptr; }
    c->scratch2_bytes = want;
    return c->scratch2;
}

// ------------------------------------------------------------------------------------------------
extern "C" {

bool tb200_link(const struct tb200_host_api *api) {
    if (api) g_host = *api;
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return false;
    if (g_host.getenv) {
        const char *v = g_host.getenv("TB200_VERBOSE");
        if (v && *v == '1') { char buf[128]; snprintf(buf, sizeof buf, "tinyblas_b200: device %d, %d SMs\n", c->device, c->sm_count); log_line(buf); }
    }
    return true;
}
int tb200_device_count(void) { std::lock_guard<std::mutex> lk(g_mu); return probe_devices(); }
int tb200_set_device(int device) {
    if (cudaSetDevice(device) != cudaSuccess) { tb200_set_error("cudaSetDevice(%d) failed", device); cudaGetLastError(); return TB200_ERR_CUDA; }
    return tb200_get_ctx() ? TB200_OK : TB200_ERR_NO_DEVICE;
}
const char *tb200_last_error(void) { return g_err; }
int tb200_sm_count(void) { tb200_ctx *c = tb200_get_ctx(); return c ? c->sm_count : 0; }
int tb200_set_stream(void *s) { tb200_ctx *c = tb200_get_ctx(); if (!c) return TB200_ERR_NO_DEVICE; c->stream = (cudaStream_t)s; return TB200_OK; }
int tb200_synchronize(void) { tb200_ctx *c = tb200_get_ctx(); if (!c) return TB200_ERR_NO_DEVICE; TB_CUDA(cudaStreamSynchronize(c->stream)); return TB200_OK; }
uint64_t tb200_launch_count(void) { return g_launches.load(); }

void *tb200_malloc(size_t bytes) { if (!tb200_get_ctx()) return nullptr; void *p = nullptr; if (cudaMalloc(&p, bytes) != cudaSuccess) { cudaGetLastError(); tb200_set_error("cudaMalloc(%zu) failed", bytes); return nullptr; } return p; }
void tb200_free(void *p) { if (p) cudaFree(p); }
int tb200_memcpy_h2d(void *dst, const void *src, size_t bytes) { tb200_ctx *c = tb200_get_ctx(); if (!c) return TB200_ERR_NO_DEVICE; TB_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c->stream)); TB_CUDA(cudaStreamSynchronize(c->stream)); return TB200_OK; }
int tb200_memcpy_d2h(void *dst, const void *src, size_t bytes) { tb200_ctx *c = tb200_get_ctx(); if (!c) return TB200_ERR_NO_DEVICE; TB_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c->stream)); TB_CUDA(cudaStreamSynchronize(c->stream)); return TB200_OK; }
int tb200_memset(void *dst, int value, size_t bytes) { tb200_ctx *c = tb200_get_ctx(); if (!c) return TB200_ERR_NO_DEVICE; TB_CUDA(cudaMemsetAsync(dst, value, bytes, c->stream)); TB_CUDA(cudaStreamSynchronize(c->stream)); return TB200_OK; }
void *tb200_host_alloc(size_t bytes) { if (!tb200_get_ctx()) return nullptr; void *p = nullptr; if (cudaMallocHost(&p, bytes) != cudaSuccess) { cudaGetLastError(); return nullptr; } return p; }
void tb200_host_free(void *p) { if (p) cudaFreeHost(p); }

// ---- resident weights --------------------------------------------------------------------------
static bool weight_geometry(int type, long rows, long k, tb200_weights *w) {
    w->type = type, w->rows = rows, w->k = k;
    switch (type) {
    case TB200_TYPE_Q4_0: if (k % QK) return false; w->qs_row_bytes = (size_t)k / QK * 16, w->sc_row_bytes = (size_t)k / QK * 2; break;
    case TB200_TYPE_Q8_0: if (k % QK) return false; w->qs_row_bytes = (size_t)k, w->sc_row_bytes = (size_t)k / QK * 2; break;
    case TB200_TYPE_Q4_K: if (k % QK_K) return false; w->qs_row_bytes = (size_t)k / 2, w->sc_row_bytes = (size_t)k / QK_K * 16; break;
    case TB200_TYPE_Q6_K: if (k % QK_K) return false; w->qs_row_bytes = (size_t)k / QK_K * 192, w->sc_row_bytes = (size_t)k / QK_K * 16, w->d_row_bytes = (size_t)k / QK_K * 2; break;
    case TB200_TYPE_F16: w->qs_row_bytes = (size_t)k * 2, w->sc_row_bytes = 0; break;
    case TB200_TYPE_Q4_1: case TB200_TYPE_Q5_0: if (k % QK) return false; w->qs_row_bytes = (size_t)k, w->sc_row_bytes = (size_t)k / QK * 4; break;
    case TB200_TYPE_Q5_K: case TB200_TYPE_IQ4_XS: if (k % QK_K) return false; w->qs_row_bytes = (size_t)k, w->sc_row_bytes = (size_t)k / QK * 4; break;
    default: return false;
    }
    return true;
}

static size_t gguf_row_bytes(int type, long k) {
    switch (type) {
    case TB200_TYPE_Q4_0: return (size_t)k / QK * 18;
    case TB200_TYPE_Q8_0: return (size_t)k / QK * 34;
    case TB200_TYPE_Q4_K: return (size_t)k / QK_K * 144;
    case TB200_TYPE_Q6_K: return (size_t)k / QK_K * 210;
    case TB200_TYPE_Q8_K: return (size_t)k / QK_K * 292;
    case TB200_TYPE_Q4_1: return (size_t)k / QK * 20;
    case TB200_TYPE_Q5_0: return (size_t)k / QK * 22;
    case TB200_TYPE_Q8_1: return (size_t)k / QK * 36;
    case TB200_TYPE_Q5_K: return (size_t)k / QK_K * 176;
    case TB200_TYPE_IQ4_XS: return (size_t)k / QK_K * 136;
    case TB200_TYPE_F16: return (size_t)k * 2;
    case TB200_TYPE_F32: return (size_t)k * 4;
    default: return 0;
    }
}

tb200_weights *tb200_weights_from_device(const void *dev_rows, int type, long rows, long k, size_t row_stride) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return nullptr;
    tb200_weights *w = new tb200_weights();
    memset(w, 0, sizeof(*w));
    if (rows < 0 || k <= 0 || !weight_geometry(type, rows, k, w)) { tb200_set_error("weights: unsupported type %d / k %ld", type, k); delete w; return nullptr; }
    if (row_stride == 0) row_stride = gguf_row_bytes(type, k);
    w->device = c->device;
    const size_t qs_bytes = round_up(w->qs_row_bytes * rows, 256), sc_bytes = round_up(w->sc_row_bytes * rows, 256),
                 d_bytes = round_up(w->d_row_bytes * rows, 256);
    w->bytes = qs_bytes + sc_bytes + d_bytes;
    void *base = nullptr;
    if (cudaMalloc(&base, w->bytes + 256) != cudaSuccess) { cudaGetLastError(); tb200_set_error("weights: cudaMalloc(%zu) failed", w->bytes); delete w; return nullptr; }
    w->qs = base;
    w->sc = sc_bytes ? (uint8_t *)base + qs_bytes : nullptr;
    w->d = d_bytes ? (uint8_t *)base + qs_bytes + sc_bytes : nullptr;
    if (tb200_launch_repack(c, dev_rows, type, rows, k, row_stride, w->qs, w->sc, w->d) != TB200_OK) { cudaFree(base); delete w; return nullptr; }
    return w;
}

tb200_weights *tb200_weights_upload(const void *host_rows, int type, long rows, long k, size_t row_stride) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return nullptr;
    const size_t rb = gguf_row_bytes(type, k);
    if (!rb) { tb200_set_error("weights: unsupported type %d", type); return nullptr; }
    if (row_stride == 0) row_stride = rb;
    const size_t total = rows ? (size_t)(rows - 1) * row_stride + rb : 0;
    void *stage = nullptr;
    if (total) {
        if (cudaMalloc(&stage, total) != cudaSuccess) { cudaGetLastError(); tb200_set_error("weights: staging cudaMalloc(%zu) failed", total); return nullptr; }
        if (cudaMemcpyAsync(stage, host_rows, total, cudaMemcpyHostToDevice, c->stream) != cudaSuccess) { cudaGetLastError(); cudaFree(stage); tb200_set_error("weights: H2D failed"); return nullptr; }
    }
    tb200_weights *w = tb200_weights_from_device(stage, type, rows, k, row_stride);
    cudaStreamSynchronize(c->stream);
    if (stage) cudaFree(stage);
    return w;
}

void tb200_weights_free(tb200_weights *w) {
    if (!w) return;
    tb200_gemm_release(w);
    if (w->qs) cudaFree(w->qs);
    delete w;
}
size_t tb200_weights_bytes(const tb200_weights *w) { return w ? w->bytes : 0; }

// ---- activation quantizers ---------------------------------------------------------------------
int tb200_quantize_q8_0_dev(const float *x, long ldx, void *y, long rows, long k) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    return tb200_launch_quantize_q8_0(c, x, ldx, rows, k, y, nullptr, nullptr);
}
int tb200_quantize_q8_K_dev(const float *x, long ldx, void *y, long rows, long k) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    return tb200_launch_quantize_q8_K(c, x, ldx, rows, k, y, nullptr, nullptr);
}

}  // extern "C"

// ---- matmul drivers ----------------------------------------------------------------------------
namespace {

struct Planar { int8_t *q; int2 *ds; __half *h; size_t bytes; };

// carve the scratch arena: planar activations (+ fp16 copy for F16 weights)
Planar carve(tb200_ctx *c, int wtype, long tokens, long k, size_t extra, void **extra_out) {
    Planar p = {nullptr, nullptr, nullptr, 0};
    size_t q_bytes = 0, ds_bytes = 0, h_bytes = 0;
    if (wtype == TB200_TYPE_F16) h_bytes = round_up((size_t)tokens * k * 2, 256);
    else q_bytes = round_up((size_t)tokens * k, 256), ds_bytes = round_up((size_t)tokens * (k / QK) * 8, 256);
    p.bytes = q_bytes + ds_bytes + h_bytes + round_up(extra, 256);
    uint8_t *base = (uint8_t *)tb200_scratch(c, p.bytes);
    if (!base) return p;
    p.q = (int8_t *)base, p.ds = (int2 *)(base + q_bytes), p.h = (__half *)(base + q_bytes + ds_bytes);
    if (extra_out) *extra_out = base + q_bytes + ds_bytes + h_bytes;
    return p;
}

// one weight matrix x prepared activations -> out[t*ldo + i]; decode kernels up to 8 tokens, tcgen05 above
int run_matmul(tb200_ctx *c, const tb200_weights *w, long row0, long nrows, const int8_t *aq, const int2 *ads, const void *af,
               int af_is_f32, long a_stride, long tokens, float *out, long ldo) {
    if (tokens > TB200_GEMV_MAX_TOKENS && tb200_gemm_available() && !af_is_f32)
        return tb200_launch_gemm(c, w, row0, nrows, aq, ads, af, a_stride, tokens, out, ldo, nullptr);
    for (long t0 = 0; t0 < tokens; t0 += TB200_GEMV_MAX_TOKENS) {
        tb200_gemv_args a;
        memset(&a, 0, sizeof(a));
        a.nmat = 1;
        a.w = w, a.row0 = row0, a.nrows = nrows;
        a.aq = aq ? aq + t0 * a_stride : nullptr;
        a.ads = ads ? ads + t0 * (a_stride / QK) : nullptr;
        a.af = af ? (const uint8_t *)af + (size_t)t0 * a_stride * (af_is_f32 ? 4 : 2) : nullptr;
        a.af_is_f32 = af_is_f32, a.a_stride = a_stride;
        a.tokens = tokens - t0 < TB200_GEMV_MAX_TOKENS ? tokens - t0 : TB200_GEMV_MAX_TOKENS;
        a.out = out + t0 * ldo, a.ldo = ldo;
        a.items = nullptr, a.n_items = 0;
        const int rc = tb200_launch_gemv(c, a);
        if (rc) return rc;
    }
    return TB200_OK;
}

}  // namespace

extern "C" {

int tb200_mul_mat(const tb200_weights *w, const float *x, long ldx, long tokens, float *out, long ldo) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    if (!w || tokens < 0 || ldx < w->k || ldo < w->rows) { tb200_set_error("mul_mat: bad arguments"); return TB200_ERR_INVALID; }
    if (tokens == 0 || w->rows == 0) return TB200_OK;
    int rc;
    if (w->type == TB200_TYPE_F16) {
        // tinyblas_cpu_sgemm.inc:121-136: B stays F32 for n <= 2, otherwise ggml converts it to F16 first (ggml.c:12345-12373)
        if (tokens <= 2) return run_matmul(c, w, 0, w->rows, nullptr, nullptr, x, 1, ldx, tokens, out, ldo);
        Planar p = carve(c, w->type, tokens, w->k, 0, nullptr);
        if (!p.h) return TB200_ERR_CUDA;
        if ((rc = tb200_launch_f32_to_f16(c, x, ldx, tokens, w->k, p.h))) return rc;
        return run_matmul(c, w, 0, w->rows, nullptr, nullptr, p.h, 0, w->k, tokens, out, ldo);
    }
    Planar p = carve(c, w->type, tokens, w->k, 0, nullptr);
    if (!p.q) return TB200_ERR_CUDA;
    if (tb200_is_kquant(w->type)) rc = tb200_launch_quantize_q8_K(c, x, ldx, tokens, w->k, nullptr, p.q, p.ds);
    else rc = tb200_launch_quantize_q8_0(c, x, ldx, tokens, w->k, nullptr, p.q, p.ds);
    if (rc) return rc;
    return run_matmul(c, w, 0, w->rows, p.q, p.ds, nullptr, 0, w->k, tokens, out, ldo);
}

int tb200_mul_mat_multi(int n, const tb200_weights *const *w, const float *x, long ldx, long tokens, float *const *out, const long *ldo) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    if (n < 1 || n > TB200_GEMV_MAX_MATS || !w || !out || !ldo || tokens < 0) { tb200_set_error("mul_mat_multi: bad arguments"); return TB200_ERR_INVALID; }
    for (int m = 0; m < n; ++m)
        if (!w[m] || w[m]->type != w[0]->type || w[m]->k != w[0]->k || ldo[m] < w[m]->rows || ldx < w[m]->k) {
            tb200_set_error("mul_mat_multi: matrices must share type and K");
            return TB200_ERR_INVALID;
        }
    if (tokens == 0) return TB200_OK;
    if ((w[0]->type != TB200_TYPE_Q4_0 && !tb200_is_kquant(w[0]->type)) || tokens > TB200_GEMV_MAX_TOKENS) {   // other types / prefill: one matmul each
        for (int m = 0; m < n; ++m) {
            const int rc = tb200_mul_mat(w[m], x, ldx, tokens, out[m], ldo[m]);
            if (rc) return rc;
        }
        return TB200_OK;
    }
    // quantize once (one launch), then ONE decode launch over the concatenated row tiles of all n matrices
    Planar p = carve(c, w[0]->type, tokens, w[0]->k, 0, nullptr);
    if (!p.q) return TB200_ERR_CUDA;
    int rc = tb200_is_kquant(w[0]->type) ? tb200_launch_quantize_q8_K(c, x, ldx, tokens, w[0]->k, nullptr, p.q, p.ds)
                                           : tb200_launch_quantize_q8_0(c, x, ldx, tokens, w[0]->k, nullptr, p.q, p.ds);
    if (rc) return rc;
    tb200_gemv_args a;
    memset(&a, 0, sizeof(a));
    a.w = w[0], a.row0 = 0, a.nrows = w[0]->rows, a.out = out[0], a.ldo = ldo[0];
    a.nmat = n;
    for (int m = 1; m < n; ++m) a.w_more[m - 1] = w[m], a.out_more[m - 1] = out[m], a.ldo_more[m - 1] = ldo[m];
    a.aq = p.q, a.ads = p.ds, a.a_stride = w[0]->k, a.tokens = tokens;
    return tb200_launch_gemv(c, a);
}

// ---- resident activations ---------------------------------------------------------------------------
struct tb200_acts {
    int wtype;          // weight type this was prepared for (decides Q8_0 / Q8_K / F16 / F32 form)
    long k, max_tokens, tokens;
    int8_t *q; int2 *ds; __half *h; const float *f32; long ldx;
    void *base;
};

tb200_acts *tb200_acts_create(int weight_type, long k, long max_tokens) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return nullptr;
    if (k <= 0 || max_tokens <= 0 || (weight_type != TB200_TYPE_F16 && k % (tb200_is_kquant(weight_type) ? QK_K : QK))) {
        tb200_set_error("acts_create: bad arguments");
        return nullptr;
    }
    tb200_acts *a = new tb200_acts();
    memset(a, 0, sizeof(*a));
    a->wtype = weight_type, a->k = k, a->max_tokens = max_tokens;
    const size_t qb = round_up((size_t)max_tokens * k, 256), dsb = round_up((size_t)max_tokens * (k / QK + 1) * 8, 256),
                 hb = round_up((size_t)max_tokens * k * 2, 256);
    const size_t total = weight_type == TB200_TYPE_F16 ? hb : qb + dsb;
    if (cudaMalloc(&a->base, total) != cudaSuccess) { cudaGetLastError(); tb200_set_error("acts_create: cudaMalloc(%zu) failed", total); delete a; return nullptr; }
    if (weight_type == TB200_TYPE_F16) a->h = (__half *)a->base;
    else a->q = (int8_t *)a->base, a->ds = (int2 *)((uint8_t *)a->base + qb);
    return a;
}

int tb200_acts_quantize(tb200_acts *a, const float *x, long ldx, long tokens) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    if (!a || tokens < 0 || tokens > a->max_tokens || ldx < a->k) { tb200_set_error("acts_quantize: bad arguments"); return TB200_ERR_INVALID; }
    a->tokens = tokens, a->f32 = nullptr, a->ldx = ldx;
    if (a->wtype == TB200_TYPE_F16) {
        if (tokens <= 2) { a->f32 = x; return TB200_OK; }   // B stays F32 for n <= 2 (tinyblas_cpu_sgemm.inc:121-136); caller keeps x alive
        return tb200_launch_f32_to_f16(c, x, ldx, tokens, a->k, a->h);
    }
    if (tb200_is_kquant(a->wtype)) return tb200_launch_quantize_q8_K(c, x, ldx, tokens, a->k, nullptr, a->q, a->ds);
    return tb200_launch_quantize_q8_0(c, x, ldx, tokens, a->k, nullptr, a->q, a->ds);
}

int tb200_mul_mat_acts(const tb200_weights *w, const tb200_acts *a, float *out, long ldo) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    if (!w || !a || a->k != w->k || ldo < w->rows) { tb200_set_error("mul_mat_acts: bad arguments"); return TB200_ERR_INVALID; }
    const bool wq = w->type != TB200_TYPE_F16, aq = a->wtype != TB200_TYPE_F16;
    if (wq != aq || (wq && tb200_is_kquant(w->type) != tb200_is_kquant(a->wtype))) {
        tb200_set_error("mul_mat_acts: activations were prepared for weight type %d, not %d", a->wtype, w->type);
        return TB200_ERR_INVALID;
    }
    if (a->tokens == 0 || w->rows == 0) return TB200_OK;
    if (!wq) {
        if (a->f32) return run_matmul(c, w, 0, w->rows, nullptr, nullptr, a->f32, 1, a->ldx, a->tokens, out, ldo);
        return run_matmul(c, w, 0, w->rows, nullptr, nullptr, a->h, 0, a->k, a->tokens, out, ldo);
    }
    return run_matmul(c, w, 0, w->rows, a->q, a->ds, nullptr, 0, a->k, a->tokens, out, ldo);
}

int tb200_mul_mat_acts_multi(int n, const tb200_weights *const *w, const tb200_acts *a, float *const *out, const long *ldo) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    if (n < 1 || n > TB200_GEMV_MAX_MATS || !w || !a || !out || !ldo) { tb200_set_error("mul_mat_acts_multi: bad arguments"); return TB200_ERR_INVALID; }
    // one launch when all matrices are Q4_0 (Q8_0 activations) or all Q4_K (Q8_K activations) and share K
    const int t0 = w[0] ? w[0]->type : -1;
    bool one_launch = (t0 == TB200_TYPE_Q4_0 || tb200_is_kquant(t0)) && a->tokens <= TB200_GEMV_MAX_TOKENS && a->wtype != TB200_TYPE_F16 &&
                      tb200_is_kquant(t0) == tb200_is_kquant(a->wtype);
    for (int m = 0; m < n && one_launch; ++m) one_launch = w[m] && w[m]->type == t0 && w[m]->k == a->k && ldo[m] >= w[m]->rows;
    if (!one_launch) {
        for (int m = 0; m < n; ++m) {
            const int rc = tb200_mul_mat_acts(w[m], a, out[m], ldo[m]);
            if (rc) return rc;
        }
        return TB200_OK;
    }
    if (a->tokens == 0) return TB200_OK;
    tb200_gemv_args g;
    memset(&g, 0, sizeof(g));
    g.w = w[0], g.row0 = 0, g.nrows = w[0]->rows, g.out = out[0], g.ldo = ldo[0], g.nmat = n;
    for (int m = 1; m < n; ++m) g.w_more[m - 1] = w[m], g.out_more[m - 1] = out[m], g.ldo_more[m - 1] = ldo[m];
    g.aq = a->q, g.ads = a->ds, g.a_stride = a->k, g.tokens = a->tokens;
    return tb200_launch_gemv(c, g);
}

// tb200_mul_mat_acts_multi with the all-gather fused into the kernel: matrix m's rows are this rank's slice starting at row
// row_lo[m] of the full output, which lives at byte offset out_offset[m] of the symmetric buffer with leading dimension
// ldo[m] (the full N). Every rank holds every slice once tb200_peers_barrier() has passed.
int tb200_mul_mat_acts_multi_gather(tb200_peers *g, int n, const tb200_weights *const *w, const tb200_acts *a, const size_t *out_offset,
                                    const long *ldo, const long *row_lo) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    if (!g || n < 1 || n > TB200_GEMV_MAX_MATS || !w || !a || !out_offset || !ldo || !row_lo) { tb200_set_error("multi_gather: bad arguments"); return TB200_ERR_INVALID; }
    const int t0 = w[0] ? w[0]->type : -1;
    for (int m = 0; m < n; ++m)
        if (!w[m] || w[m]->type != t0 || w[m]->k != a->k || row_lo[m] < 0 || ldo[m] < row_lo[m] + w[m]->rows || (out_offset[m] & 3)) {
            tb200_set_error("multi_gather: matrix %d does not fit its slot", m);
            return TB200_ERR_INVALID;
        }
    if (a->tokens > TB200_GEMV_MAX_TOKENS) {
        // prefill: one tcgen05 launch per matrix, the epilogue stores each 128-byte output segment into every rank's buffer
        if (!tb200_gemm_available()) return TB200_ERR_UNSUPPORTED;
        const bool wq = t0 != TB200_TYPE_F16, aq_ok = a->wtype != TB200_TYPE_F16;
        if (wq != aq_ok || (wq && tb200_is_kquant(t0) != tb200_is_kquant(a->wtype)) || (!wq && a->f32)) return TB200_ERR_UNSUPPORTED;
        char *base = (char *)g->out_base[g->rank];
        for (int m = 0; m < n; ++m) {
            float *out = (float *)(base + out_offset[m]) + row_lo[m];
            const int rc = tb200_launch_gemm(c, w[m], 0, w[m]->rows, wq ? a->q : nullptr, wq ? a->ds : nullptr, wq ? nullptr : (const void *)a->h, a->k,
                                             a->tokens, out, ldo[m], g);
            if (rc) return rc;
        }
        return TB200_OK;
    }
    if ((t0 != TB200_TYPE_Q4_0 && !tb200_is_kquant(t0)) || tb200_is_kquant(t0) != tb200_is_kquant(a->wtype) || a->wtype == TB200_TYPE_F16)
        return TB200_ERR_UNSUPPORTED;
    if (a->tokens == 0) return TB200_OK;
    char *base = (char *)g->out_base[g->rank];
    tb200_gemv_args ga;
    memset(&ga, 0, sizeof(ga));
    ga.nmat = n;
    ga.w = w[0], ga.row0 = 0, ga.nrows = w[0]->rows, ga.out = (float *)(base + out_offset[0]) + row_lo[0], ga.ldo = ldo[0];
    for (int m = 1; m < n; ++m) ga.w_more[m - 1] = w[m], ga.out_more[m - 1] = (float *)(base + out_offset[m]) + row_lo[m], ga.ldo_more[m - 1] = ldo[m];
    ga.aq = a->q, ga.ads = a->ds, ga.a_stride = a->k, ga.tokens = a->tokens;
    ga.world = g->world, ga.local_base = base;
    for (int p = 0; p < g->world; ++p) ga.peer_base[p] = g->out_base[p];
    return tb200_launch_gemv(c, ga);
}

void tb200_acts_free(tb200_acts *a) {
    if (!a) return;
    if (a->base) cudaFree(a->base);
    delete a;
}

// which (Atype, Btype, n) the reference's x86 build services (tinyblas_cpu_sgemm.inc:46-240, 274-331),
// restricted to the weight types this module holds (F16, Q8_0, Q4_0, Q4_K)
static bool serviced(int Atype, int Btype, int Ctype, long n) {
    if (Ctype != TB200_TYPE_F32) return false;
    switch (Atype) {
    case TB200_TYPE_Q4_0: case TB200_TYPE_Q8_0: case TB200_TYPE_Q5_0: return Btype == TB200_TYPE_Q8_0;
    case TB200_TYPE_Q4_1: return Btype == TB200_TYPE_Q8_1;   // iqk mul_mat_qX_1_q8_1_T<Q4_1_Unpacker> (iqk_mul_mat.inc:1442)
    case TB200_TYPE_Q4_K: case TB200_TYPE_Q6_K: case TB200_TYPE_Q5_K: case TB200_TYPE_IQ4_XS: return Btype == TB200_TYPE_Q8_K;
    case TB200_TYPE_F16: return Btype == TB200_TYPE_F16 || (Btype == TB200_TYPE_F32 && n <= 2);
    default: return false;
    }
}

int tb200_sgemm_dev(const tb200_weights *w, long n, const void *B, long ldb, int Btype, float *C, long ldc) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    if (!w || n < 0 || ldc < w->rows) { tb200_set_error("sgemm_dev: bad arguments"); return TB200_ERR_INVALID; }
    if (!serviced(w->type, Btype, TB200_TYPE_F32, n)) return TB200_ERR_UNSUPPORTED;
    if (n == 0 || w->rows == 0) return TB200_OK;
    if (w->type == TB200_TYPE_F16) {
        if (ldb < w->k) { tb200_set_error("sgemm_dev: ldb < k"); return TB200_ERR_INVALID; }
        return run_matmul(c, w, 0, w->rows, nullptr, nullptr, B, Btype == TB200_TYPE_F32, ldb, n, C, ldc);
    }
    const long blck = tb200_is_kquant(w->type) ? QK_K : QK;
    if (ldb < w->k / blck) { tb200_set_error("sgemm_dev: ldb < k"); return TB200_ERR_INVALID; }
    Planar p = carve(c, w->type, n, w->k, 0, nullptr);
    if (!p.q) return TB200_ERR_CUDA;
    int rc = tb200_is_kquant(w->type) ? tb200_launch_unpack_q8_K(c, B, ldb, n, w->k, p.q, p.ds)
             : Btype == TB200_TYPE_Q8_1 ? tb200_launch_unpack_q8_1(c, B, ldb, n, w->k, p.q, p.ds)
                                        : tb200_launch_unpack_q8_0(c, B, ldb, n, w->k, p.q, p.ds);
    if (rc) return rc;
    return run_matmul(c, w, 0, w->rows, p.q, p.ds, nullptr, 0, w->k, n, C, ldc);
}

int tb200_blockdot_dev(const tb200_weights *w, const void *B, int Btype, int32_t *out) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    if (!w || w->type == TB200_TYPE_F16 || !serviced(w->type, Btype, TB200_TYPE_F32, 1)) return TB200_ERR_UNSUPPORTED;
    Planar p = carve(c, w->type, 1, w->k, 0, nullptr);
    if (!p.q) return TB200_ERR_CUDA;
    const long blck = tb200_is_kquant(w->type) ? QK_K : QK;
    int rc = tb200_is_kquant(w->type) ? tb200_launch_unpack_q8_K(c, B, w->k / blck, 1, w->k, p.q, p.ds)
                                       : tb200_launch_unpack_q8_0(c, B, w->k / blck, 1, w->k, p.q, p.ds);
    if (rc) return rc;
    return tb200_launch_blockdot(c, w, p.q, p.ds, out);
}

}  // extern "C"

// exported for tb200_host.cu (host-pointer shims) and tb200_mixmul.cu
int tb200_run_matmul(tb200_ctx *c, const tb200_weights *w, long row0, long nrows, const int8_t *aq, const int2 *ads, const void *af,
                     int af_is_f32, long a_stride, long tokens, float *out, long ldo) {
    return run_matmul(c, w, row0, nrows, aq, ads, af, af_is_f32, a_stride, tokens, out, ldo);
}
bool tb200_serviced(int Atype, int Btype, int Ctype, long n) { return serviced(Atype, Btype, Ctype, n); }
size_t tb200_gguf_row_bytes(int type, long k) { return gguf_row_bytes(type, k); }
