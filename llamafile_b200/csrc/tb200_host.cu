// The reference's own entry points on HOST pointers (include/tinyblas_b200.h part 1):
//   llamafile_sgemm, llamafile_mixmul, llamafile_mixmul_needs, llamafile_mixmul_iqk  (llamafile/sgemm.h:23-28, 76-77)
//   quantize_row_q8_0 / quantize_row_q8_K / ggml_fp32_to_fp16_row as tb200_* (from_float of the vec_dot types)
// Same argument meaning and true/false behaviour as the CPU implementation; the work itself runs on the
// GPU: weights are uploaded once and cached by host address, activations go H2D, results come D2H, and the
// call returns when the result is in the caller's buffer. Nothing here computes on the CPU.
#include <string.h>

#include <map>
#include <mutex>
#include <tuple>

#include "tb200_common.cuh"

int tb200_run_matmul(tb200_ctx *c, const tb200_weights *w, long row0, long nrows, const int8_t *aq, const int2 *ads, const void *af,
                     int af_is_f32, long a_stride, long tokens, float *out, long ldo);
bool tb200_serviced(int Atype, int Btype, int Ctype, long n);
size_t tb200_gguf_row_bytes(int type, long k);

namespace {

std::mutex g_cache_mu;
std::map<std::tuple<const void *, long, long, size_t, int>, tb200_weights *> g_cache;

// device staging for host-pointer calls, separate from the kernels' scratch arena
struct Io { void *p = nullptr; size_t cap = 0; };
Io g_io[16];

void *io_buffer(tb200_ctx *c, size_t bytes) {
    Io &io = g_io[c->device];
    if (bytes <= io.cap) return io.p;
    cudaStreamSynchronize(c->stream);
    if (io.p) cudaFree(io.p);
    io.p = nullptr, io.cap = 0;
    const size_t want = (bytes + bytes / 2 + (1 << 20) - 1) >> 20 << 20;
    if (cudaMalloc(&io.p, want) != cudaSuccess) { cudaGetLastError(); tb200_set_error("io: cudaMalloc(%zu) failed", want); return nullptr; }
    io.cap = want;
    return io.p;
}

tb200_weights *cached_weights(const void *A, int type, long rows, long k, size_t row_stride) {
    std::lock_guard<std::mutex> lk(g_cache_mu);
    const auto key = std::make_tuple(A, rows, k, row_stride, type);
    auto it = g_cache.find(key);
    if (it != g_cache.end()) return it->second;
    tb200_weights *w = tb200_weights_upload(A, type, rows, k, row_stride);
    if (w) g_cache[key] = w;
    return w;
}

int blck(int type) { return (type == TB200_TYPE_Q4_K || type == TB200_TYPE_Q8_K) ? QK_K : (type == TB200_TYPE_Q4_0 || type == TB200_TYPE_Q8_0) ? QK : 1; }
size_t tsize(int type) { return tb200_gguf_row_bytes(type, blck(type)); }
size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace

extern "C" {

TB200_API void tb200_weights_cache_clear(void) {
    std::lock_guard<std::mutex> lk(g_cache_mu);
    for (auto &kv : g_cache) tb200_weights_free(kv.second);
    g_cache.clear();
}

bool llamafile_sgemm(long m, long n, long k, const void *A, long lda, const void *B, long ldb, void *C, long ldc, int ith, int nth,
                     int Atype, int Btype, int Ctype) {
    // the reference asserts these (tinyblas_cpu_sgemm.inc:277-284); a violated precondition is "not handled"
    if (m < 0 || n < 0 || k < 0 || lda < k || ldb < k || ldc < m || nth <= 0 || ith < 0 || ith >= nth) return false;
    if (!tb200_serviced(Atype, Btype, Ctype, n)) return false;
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return false;
    if (ith != 0) return true;   // thread 0 owns the whole job; see the threading contract in the header
    if (m == 0 || n == 0) return true;
    const long ke = k * blck(Atype);
    // iqk (Q4_0, Q4_K) takes rows ggml_row_size apart and ignores lda/ldb (iqk_mul_mat.inc:191-199); tinyBLAS honours them
    const bool iqk = Atype == TB200_TYPE_Q4_0 || Atype == TB200_TYPE_Q4_K;
    const size_t a_stride = iqk ? tb200_gguf_row_bytes(Atype, ke) : (size_t)lda * tsize(Atype);
    const long ldb_eff = iqk ? k : ldb;
    tb200_weights *w = cached_weights(A, Atype, m, ke, a_stride);
    if (!w) return false;
    const size_t b_bytes = ((size_t)(n - 1) * ldb_eff + k) * tsize(Btype), c_bytes = (size_t)n * m * 4;
    uint8_t *io = (uint8_t *)io_buffer(c, up256(b_bytes) + c_bytes);
    if (!io) return false;
    float *dC = (float *)(io + up256(b_bytes));
    if (cudaMemcpyAsync(io, B, b_bytes, cudaMemcpyHostToDevice, c->stream) != cudaSuccess) { cudaGetLastError(); return false; }
    if (tb200_sgemm_dev(w, n, io, ldb_eff, Btype, dC, m) != TB200_OK) return false;   // ldb: elements for F16/F32, blocks for quants
    if (cudaMemcpy2DAsync(C, (size_t)ldc * 4, dC, (size_t)m * 4, (size_t)m * 4, (size_t)n, cudaMemcpyDeviceToHost, c->stream) != cudaSuccess) { cudaGetLastError(); return false; }
    return cudaStreamSynchronize(c->stream) == cudaSuccess;
}

// llamafile/tinyblas_cpu_mixmul.inc:97-111 + tinyblas_cpu_mixmul_amd_avx.cpp:8-18 (arena arithmetic only)
size_t llamafile_mixmul_needs(const struct ggml_tensor *weights_, const struct ggml_tensor *thought_, const struct ggml_tensor *plan_) {
    const tb200_ggml_tensor_view *w = (const tb200_ggml_tensor_view *)weights_, *x = (const tb200_ggml_tensor_view *)thought_,
                                 *p = (const tb200_ggml_tensor_view *)plan_;
    const size_t cols = w->ne[0], experts = w->ne[2], thinkers = p->ne[0], tasks = x->ne[1], tokens = x->ne[2];
    const size_t ldq = (cols * 2 + 63) & ~(size_t)63;
    const size_t wsize = 0x7ffff000, wdata = 0x1000, wdata_aligned = (wdata + 4095) & ~(size_t)4095;
    size_t allocated = 0;
    auto alloc = [&](size_t align, size_t need) {
        const size_t base = (allocated + align - 1) & ~(align - 1), toto = base + need;
        if (toto >= allocated && toto <= wsize) { allocated = toto; return true; }
        return false;
    };
    if (!alloc(4096, tokens * tasks * ldq) || !alloc(64, 8 * experts * tokens * thinkers) || !alloc(64, 8 * experts * tokens * thinkers) ||
        !alloc(8, 8 * experts))
        return 0;
    return (wdata_aligned - wdata) + allocated;
}

bool llamafile_mixmul(const struct ggml_compute_params *params_, const struct ggml_tensor *weights_, const struct ggml_tensor *thought_,
                      const struct ggml_tensor *plan_, struct ggml_tensor *result_) {
    const tb200_ggml_compute_params_view *params = (const tb200_ggml_compute_params_view *)params_;
    const tb200_ggml_tensor_view *w = (const tb200_ggml_tensor_view *)weights_, *x = (const tb200_ggml_tensor_view *)thought_,
                                 *pl = (const tb200_ggml_tensor_view *)plan_;
    tb200_ggml_tensor_view *out = (tb200_ggml_tensor_view *)result_;
    const long rows = w->ne[1], cols = w->ne[0], experts = w->ne[2], thinkers = pl->ne[0], tasks = x->ne[1], tokens = x->ne[2];
    // allocate_shared_memory() must succeed in the caller's wdata, as in the reference (:97-107, :390-398):
    // offsets are measured from the 4096-aligned start and compared with params->wsize
    {
        const size_t need = llamafile_mixmul_needs(weights_, thought_, plan_);
        if (!need || need > params->wsize) return false;
    }
    // invariants the reference asserts (:116-135)
    if (!(tasks <= thinkers && thinkers <= experts && tokens == pl->ne[1] && rows == out->ne[0] && cols == x->ne[0] &&
          tokens == out->ne[2] && thinkers == out->ne[1] && pl->type == TB200_TYPE_I32 && params->nth > 0 && params->ith < params->nth))
        return false;
    if (pl->ne[2] != 1 || pl->ne[3] != 1 || out->ne[3] != 1 || w->ne[3] != 1 || x->ne[3] != 1) return false;
    const size_t wts = tsize(w->type);
    if (!wts || w->nb[1] % wts) return false;                                              // :138
    if (out->nb[0] != 4 || x->nb[0] != tsize(x->type) || w->nb[0] != wts) return false;    // :142-147
    if (out->type != TB200_TYPE_F32) return false;
    if (w->type != TB200_TYPE_Q4_0 && w->type != TB200_TYPE_Q8_0 && w->type != TB200_TYPE_F16) return false;
    if (x->type != TB200_TYPE_F32) return false;   // pre-converted thought (Q8_0/F16) is never produced by ggml for this op
    if (cols % (w->type == TB200_TYPE_F16 ? 16 : 32)) return false;                        // :273
    if (w->nb[2] != (size_t)rows * w->nb[1]) return false;                                 // experts must be stacked
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return false;
    if (params->ith != 0) return true;
    if (tokens == 0 || rows == 0) return true;
    tb200_weights *dw = cached_weights(w->data, w->type, rows * experts, cols, w->nb[1]);
    if (!dw) return false;
    const size_t x_bytes = (size_t)tokens * tasks * cols * 4, p_bytes = (size_t)tokens * thinkers * 4, o_bytes = (size_t)tokens * thinkers * rows * 4;
    uint8_t *io = (uint8_t *)io_buffer(c, up256(x_bytes) + up256(p_bytes) + o_bytes);
    if (!io) return false;
    float *dx = (float *)io;
    int32_t *dp = (int32_t *)(io + up256(x_bytes));
    float *dout = (float *)(io + up256(x_bytes) + up256(p_bytes));
    bool ok = true;
    // gather the (possibly strided) host tensors into contiguous device buffers
    for (long t = 0; t < tokens && ok; ++t)
        ok = cudaMemcpy2DAsync(dx + (size_t)t * tasks * cols, (size_t)cols * 4, (const char *)x->data + t * x->nb[2], x->nb[1], (size_t)cols * 4,
                               (size_t)tasks, cudaMemcpyHostToDevice, c->stream) == cudaSuccess;
    bool all_valid = true;
    if (ok) {
        if (pl->nb[0] == 4) {
            ok = cudaMemcpy2DAsync(dp, (size_t)thinkers * 4, pl->data, pl->nb[1], (size_t)thinkers * 4, (size_t)tokens, cudaMemcpyHostToDevice, c->stream) == cudaSuccess;
        } else ok = false;
        for (long t = 0; t < tokens; ++t)
            for (long th = 0; th < thinkers; ++th) {
                const int32_t e = *(const int32_t *)((const char *)pl->data + t * pl->nb[1] + th * pl->nb[0]);
                if (e < 0 || e >= experts) all_valid = false;
            }
    }
    // rows naming no expert stay untouched in the reference: carry the caller's bytes through the device copy
    if (ok && !all_valid)
        for (long t = 0; t < tokens && ok; ++t)
            ok = cudaMemcpy2DAsync(dout + (size_t)t * thinkers * rows, (size_t)rows * 4, (const char *)out->data + t * out->nb[2], out->nb[1],
                                   (size_t)rows * 4, (size_t)thinkers, cudaMemcpyHostToDevice, c->stream) == cudaSuccess;
    if (ok) ok = tb200_mixmul_dev(dw, (int)experts, dx, (int)tasks, dp, (int)thinkers, tokens, dout) == TB200_OK;
    for (long t = 0; t < tokens && ok; ++t)
        ok = cudaMemcpy2DAsync((char *)out->data + t * out->nb[2], out->nb[1], dout + (size_t)t * thinkers * rows, (size_t)rows * 4,
                               (size_t)rows * 4, (size_t)thinkers, cudaMemcpyDeviceToHost, c->stream) == cudaSuccess;
    if (!ok) cudaGetLastError();
    return cudaStreamSynchronize(c->stream) == cudaSuccess && ok;
}

// iqk_mul_mat_moe (llamafile/iqk_mul_mat.inc:204-221): A = one expert's Nx rows, B = all quantized rows
// [ne12][ne11], C row for mapping y = C + i1*nb1 + i2*nb2, B row = (i1 % ne11) + i2*ne11.
bool llamafile_mixmul_iqk(long Nx, long Ny, long ne00, int ne11, int typeA, const void *A, const void *B, float *C, long nb1, long nb2,
                          const void *vrow_mapping, int ith, int nth) {
    if (typeA != TB200_TYPE_Q4_K && typeA != TB200_TYPE_Q4_0) return false;   // the other iqk types are outside this module
    if (!vrow_mapping || Nx < 0 || Ny < 0 || ne11 <= 0 || nth <= 0 || ith < 0 || ith >= nth || ne00 % blck(typeA)) return false;
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return false;
    if (ith != 0) return true;
    if (Nx == 0 || Ny == 0) return true;
    const int Btype = typeA == TB200_TYPE_Q4_K ? TB200_TYPE_Q8_K : TB200_TYPE_Q8_0;
    tb200_weights *w = cached_weights(A, typeA, Nx, ne00, 0);
    if (!w) return false;
    const int32_t *rm = (const int32_t *)vrow_mapping;
    const size_t row_b = tb200_gguf_row_bytes(Btype, ne00);
    // gather the Ny mapped B rows into a contiguous staging block, run one dense sgemm, scatter rows of C back
    uint8_t *io = (uint8_t *)io_buffer(c, up256(row_b * Ny) + (size_t)Ny * Nx * 4);
    if (!io) return false;
    float *dC = (float *)(io + up256(row_b * Ny));
    bool ok = true;
    for (long y = 0; y < Ny && ok; ++y) {
        const long brow = (long)(rm[2 * y] % ne11) + (long)rm[2 * y + 1] * ne11;
        ok = cudaMemcpyAsync(io + y * row_b, (const char *)B + brow * row_b, row_b, cudaMemcpyHostToDevice, c->stream) == cudaSuccess;
    }
    if (ok) ok = tb200_sgemm_dev(w, Ny, io, ne00 / blck(typeA), Btype, dC, Nx) == TB200_OK;
    for (long y = 0; y < Ny && ok; ++y) {
        float *dst = C + (long)rm[2 * y] * (nb1 / 4) + (long)rm[2 * y + 1] * (nb2 / 4);
        ok = cudaMemcpyAsync(dst, dC + y * Nx, (size_t)Nx * 4, cudaMemcpyDeviceToHost, c->stream) == cudaSuccess;
    }
    if (!ok) cudaGetLastError();
    return cudaStreamSynchronize(c->stream) == cudaSuccess && ok;
}

static bool quantize_row_host(const float *x, void *y, int64_t k, int type) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c || k < 0) return false;
    if (k == 0) return true;
    const size_t out_b = type == TB200_TYPE_F16 ? (size_t)k * 2 : tb200_gguf_row_bytes(type, k);
    if (!out_b) return false;
    uint8_t *io = (uint8_t *)io_buffer(c, up256((size_t)k * 4) + out_b);
    if (!io) return false;
    uint8_t *dy = io + up256((size_t)k * 4);
    if (cudaMemcpyAsync(io, x, (size_t)k * 4, cudaMemcpyHostToDevice, c->stream) != cudaSuccess) { cudaGetLastError(); return false; }
    int rc;
    if (type == TB200_TYPE_Q8_0) rc = tb200_launch_quantize_q8_0(c, (const float *)io, k, 1, k, dy, nullptr, nullptr);
    else if (type == TB200_TYPE_Q8_K) rc = tb200_launch_quantize_q8_K(c, (const float *)io, k, 1, k, dy, nullptr, nullptr);
    else rc = tb200_launch_f32_to_f16(c, (const float *)io, k, 1, k, (__half *)dy);
    if (rc) return false;
    if (cudaMemcpyAsync(y, dy, out_b, cudaMemcpyDeviceToHost, c->stream) != cudaSuccess) { cudaGetLastError(); return false; }
    return cudaStreamSynchronize(c->stream) == cudaSuccess;
}
bool tb200_quantize_row_q8_0(const float *x, void *y, int64_t k) { return k % QK == 0 && quantize_row_host(x, y, k, TB200_TYPE_Q8_0); }
bool tb200_quantize_row_q8_K(const float *x, void *y, int64_t k) { return k % QK_K == 0 && quantize_row_host(x, y, k, TB200_TYPE_Q8_K); }
bool tb200_fp32_to_fp16_row(const float *x, uint16_t *y, int64_t k) { return quantize_row_host(x, y, k, TB200_TYPE_F16); }

}  // extern "C"
