// tb200_mul_mat_noalloc: the decode-size tb200_mul_mat for callers that may not allocate (CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc). The quantized activations that tb200_mul_mat keeps in the context's growable
// cudaMalloc arena live here in a file-scope device array of fixed size, so the call path contains no cudaMalloc, cudaFree or
// allocating library routine: one activation-quantizer launch into the array, one decode launch reading it (the same kernels
// and launchers tb200_mul_mat uses, so the results are bit-identical).
//
// The array exists once per GPU that loads the library (TB200_NOALLOC_SCRATCH_BYTES of HBM held until exit) and is shared by
// every call on that GPU: it keeps the previous call's contents, and two calls that overlap on the GPU (different streams, or
// graph branches) race on it. Nothing in it has to start at zero: the quantizer rewrites every byte the decode kernel reads
// (q[tokens][k] and one {d, sum} record per 32-block) before the kernel starts, in stream order.
#include <mutex>

#include "tb200_common.cuh"

static __device__ __align__(256) unsigned char g_noalloc_scratch[TB200_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 16;
std::mutex g_mu;
unsigned char *g_scratch_ptr[kMaxDevices];   // device address of g_noalloc_scratch on each GPU (the symbol itself is a host-side stand-in)

size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Once per device, with that device current (c is the current device's context): resolve the array's device address.
unsigned char *scratch_base(tb200_ctx *c) {
    if (c->device < 0 || c->device >= kMaxDevices) { tb200_set_error("mul_mat_noalloc: device %d out of range", c->device); return nullptr; }
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_scratch_ptr[c->device]) {
        unsigned char *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            tb200_set_error("mul_mat_noalloc: cudaGetSymbolAddress failed on device %d: %s", c->device, cudaGetErrorString(e));
            return nullptr;
        }
        g_scratch_ptr[c->device] = p;
    }
    return g_scratch_ptr[c->device];
}

}  // namespace

extern "C" int tb200_mul_mat_noalloc(const tb200_weights *w, const float *x, long ldx, long tokens, float *out, long ldo) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    if (!w || !x || !out || tokens < 0 || ldx < w->k || ldo < w->rows) { tb200_set_error("mul_mat_noalloc: bad arguments"); return TB200_ERR_INVALID; }
    if (w->device != c->device) { tb200_set_error("mul_mat_noalloc: weights live on device %d, current device is %d", w->device, c->device); return TB200_ERR_INVALID; }
    if (tokens > TB200_GEMV_MAX_TOKENS) {
        tb200_set_error("mul_mat_noalloc: %ld tokens exceed the limit of %d (decode-size calls only; use tb200_mul_mat)", tokens, TB200_GEMV_MAX_TOKENS);
        return TB200_ERR_INVALID;
    }
    // planar activations: int8 q[tokens][k], then {float d, int32 sum}[tokens][k/32] (F16 weights, 3..8 tokens: fp16 x[tokens][k])
    const bool f16 = w->type == TB200_TYPE_F16;
    const size_t q_bytes = f16 ? 0 : round_up((size_t)tokens * w->k, 256), ds_bytes = f16 ? 0 : round_up((size_t)tokens * (w->k / QK) * 8, 256),
                 h_bytes = f16 && tokens > 2 ? round_up((size_t)tokens * w->k * 2, 256) : 0;
    if (q_bytes + ds_bytes + h_bytes > (size_t)TB200_NOALLOC_SCRATCH_BYTES) {
        tb200_set_error("mul_mat_noalloc: %ld tokens x k=%ld need %zu bytes of activation scratch, the fixed array holds %d", tokens, w->k,
                        q_bytes + ds_bytes + h_bytes, TB200_NOALLOC_SCRATCH_BYTES);
        return TB200_ERR_INVALID;
    }
    if (tokens == 0 || w->rows == 0) return TB200_OK;
    unsigned char *base = scratch_base(c);
    if (!base) return TB200_ERR_CUDA;

    tb200_gemv_args a = {};
    a.nmat = 1;
    a.w = w, a.row0 = 0, a.nrows = w->rows;
    a.tokens = tokens, a.out = out, a.ldo = ldo;
    int rc;
    if (f16) {
        // tinyblas_cpu_sgemm.inc:121-136: B stays F32 for n <= 2, otherwise it is converted to F16 first (ggml.c:12345-12373)
        if (tokens <= 2) a.af = x, a.af_is_f32 = 1, a.a_stride = ldx;
        else {
            if ((rc = tb200_launch_f32_to_f16(c, x, ldx, tokens, w->k, (__half *)base))) return rc;
            a.af = base, a.af_is_f32 = 0, a.a_stride = w->k;
        }
    } else {
        int8_t *q = (int8_t *)base;
        int2 *ds = (int2 *)(base + q_bytes);
        rc = tb200_is_kquant(w->type) ? tb200_launch_quantize_q8_K(c, x, ldx, tokens, w->k, nullptr, q, ds)
                                      : tb200_launch_quantize_q8_0(c, x, ldx, tokens, w->k, nullptr, q, ds);
        if (rc) return rc;
        a.aq = q, a.ads = ds, a.a_stride = w->k;
    }
    return tb200_launch_gemv(c, a);
}
