// The matmul's neighbours in the reference graph for ANY token count. The decode plan (tb200_plan.cu) fuses
// them into its persistent kernel for one token; these are the batch forms that feed the resident-activation API:
//   llm_build_norm : ggml_rms_norm (ggml.c:11850-11896) then ggml_mul by the norm weight      -> from_float of the vec_dot type
//   llm_build_ffn  : ggml_silu(gate) (ggml.c:2393) then ggml_mul by up (LLM_FFN_SILU, PAR)    -> from_float
//   MoE tail       : ggml_mul(experts, weights) then ggml_add over the used experts           (llama.cpp/src/llama.cpp:8303-8318)
// Each transform kernel writes the fp32 values the reference graph would hold at that point (same IEEE operations in the same
// order, explicit _rn intrinsics), and the module's bit-exact quantizer (tb200_quantize.cu) turns them into activations: the
// result equals "reference ops, then quantize_row_q8_0 / q8_K" wherever the fp32 values agree bit for bit (RMSNorm: always, up
// to the last bit of a double-precision sum; SiLU: within a few ulp of ggml's SIMD polynomial expf, see oracle_silu_mul).
#include "tb200_common.cuh"

struct tb200_acts;   // tb200_api.cu
int tb200_acts_quantize(tb200_acts *a, const float *x, long ldx, long tokens);

namespace {

// one CTA per token row
__global__ void __launch_bounds__(256) rmsnorm_mul_kernel(const float *__restrict__ x, long ldx, const float *__restrict__ w, float eps, long k,
                                                          float *__restrict__ y, long ldy) {
    pdl_launch_dependents();
    pdl_wait();
    const float *row = x + (long)blockIdx.x * ldx;
    float *out = y + (long)blockIdx.x * ldy;
    double s = 0.0;
    for (long i = threadIdx.x; i < k; i += blockDim.x) s += (double)__fmul_rn(row[i], row[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    __shared__ double s_part[8];
    __shared__ float s_scale;
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < 8; ++i) t += s_part[i];
        const float mean = (float)(t / (double)k);
        s_scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
    }
    __syncthreads();
    const float scale = s_scale;
    if (w) for (long i = threadIdx.x; i < k; i += blockDim.x) out[i] = __fmul_rn(__fmul_rn(row[i], scale), w[i]);
    else for (long i = threadIdx.x; i < k; i += blockDim.x) out[i] = __fmul_rn(row[i], scale);   // ggml_rms_norm alone (ggml_vec_scale_f32)
}

__global__ void __launch_bounds__(256) swiglu_kernel(const float *__restrict__ gate, const float *__restrict__ up, long ld, long rows, long k,
                                                     float *__restrict__ y, long ldy) {
    pdl_launch_dependents();
    pdl_wait();
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i >= rows * k) return;
    const long r = i / k, c = i % k;
    const float g = gate[r * ld + c];
    y[r * ldy + c] = __fmul_rn(__fdiv_rn(g, __fadd_rn(1.0f, expf(-g))), up[r * ld + c]);
}

// out[t][r] = ((e[t][0][r]*w[t][0]) + e[t][1][r]*w[t][1]) + ... : products rounded, then added in expert order
__global__ void __launch_bounds__(256) moe_combine_kernel(const float *__restrict__ e, const float *__restrict__ w, long tokens, int used, long rows,
                                                          float *__restrict__ out, long ldo) {
    pdl_launch_dependents();
    pdl_wait();
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i >= tokens * rows) return;
    const long t = i / rows, r = i % rows;
    const float *et = e + t * used * rows + r;
    float acc = __fmul_rn(et[0], w[t * used]);
    for (int j = 1; j < used; ++j) acc = __fadd_rn(acc, __fmul_rn(et[(long)j * rows], w[t * used + j]));
    out[t * ldo + r] = acc;
}

// the graph's plain elementwise ops, for a ggml backend that keeps the matmul's neighbours on the device when they cannot be fused
// (more than one token, or an intermediate somebody else reads): y = a (op) b with b either a full [tokens][k] tensor or one row
// broadcast over the tokens (ggml_mul by a norm weight), y = silu(x)
enum { kOpMul = 0, kOpAdd = 1, kOpSilu = 2 };
template <int kOp>
__global__ void __launch_bounds__(256) elementwise_kernel(const float *__restrict__ a, long lda, const float *__restrict__ b, long ldb, long tokens, long k,
                                                          float *__restrict__ y, long ldy) {
    pdl_launch_dependents();
    pdl_wait();
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i >= tokens * k) return;
    const long r = i / k, c = i % k;
    const float av = a[r * lda + c];
    float out;
    if (kOp == kOpSilu) out = __fdiv_rn(av, __fadd_rn(1.0f, expf(-av)));
    else if (kOp == kOpMul) out = __fmul_rn(av, b[r * ldb + c]);
    else out = __fadd_rn(av, b[r * ldb + c]);
    y[r * ldy + c] = out;
}

}  // namespace

extern "C" {

int tb200_rms_norm_dev(const float *x, long ldx, float eps, long tokens, long k, float *y, long ldy) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    if (!x || !y || tokens < 0 || k <= 0 || ldx < k || ldy < k) { tb200_set_error("rms_norm: bad arguments"); return TB200_ERR_INVALID; }
    if (tokens == 0) return TB200_OK;
    TB_CUDA(tb200_launch(rmsnorm_mul_kernel, dim3((unsigned)tokens), dim3(256), 0, c->stream, x, ldx, (const float *)nullptr, eps, k, y, ldy));
    return TB200_OK;
}

static int elementwise(int op, const float *a, long lda, const float *b, long ldb, long tokens, long k, float *y, long ldy) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    if (!a || !y || (op != kOpSilu && !b) || tokens < 0 || k <= 0 || lda < k || ldy < k || (ldb != 0 && ldb < k)) { tb200_set_error("elementwise: bad arguments"); return TB200_ERR_INVALID; }
    if (tokens == 0) return TB200_OK;
    const dim3 grid((unsigned)((tokens * k + 255) / 256));
    if (op == kOpMul) TB_CUDA(tb200_launch(elementwise_kernel<kOpMul>, grid, dim3(256), 0, c->stream, a, lda, b, ldb, tokens, k, y, ldy));
    else if (op == kOpAdd) TB_CUDA(tb200_launch(elementwise_kernel<kOpAdd>, grid, dim3(256), 0, c->stream, a, lda, b, ldb, tokens, k, y, ldy));
    else TB_CUDA(tb200_launch(elementwise_kernel<kOpSilu>, grid, dim3(256), 0, c->stream, a, lda, b, ldb, tokens, k, y, ldy));
    return TB200_OK;
}
int tb200_mul_dev(const float *a, long lda, const float *b, long ldb, long tokens, long k, float *y, long ldy) { return elementwise(kOpMul, a, lda, b, ldb, tokens, k, y, ldy); }
int tb200_add_dev(const float *a, long lda, const float *b, long ldb, long tokens, long k, float *y, long ldy) { return elementwise(kOpAdd, a, lda, b, ldb, tokens, k, y, ldy); }
int tb200_silu_dev(const float *x, long ldx, long tokens, long k, float *y, long ldy) { return elementwise(kOpSilu, x, ldx, nullptr, 0, tokens, k, y, ldy); }

int tb200_rms_norm_mul_dev(const float *x, long ldx, const float *norm_weight, float eps, long tokens, long k, float *y, long ldy) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    if (!x || !norm_weight || !y || tokens < 0 || k <= 0 || ldx < k || ldy < k || !(eps > 0.f)) { tb200_set_error("rms_norm_mul: bad arguments"); return TB200_ERR_INVALID; }
    if (tokens == 0) return TB200_OK;
    TB_CUDA(tb200_launch(rmsnorm_mul_kernel, dim3((unsigned)tokens), dim3(256), 0, c->stream, x, ldx, norm_weight, eps, k, y, ldy));
    return TB200_OK;
}

int tb200_silu_mul_dev(const float *gate, const float *up, long ld, long tokens, long k, float *y, long ldy) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    if (!gate || !up || !y || tokens < 0 || k <= 0 || ld < k || ldy < k) { tb200_set_error("silu_mul: bad arguments"); return TB200_ERR_INVALID; }
    if (tokens == 0) return TB200_OK;
    const long n = tokens * k;
    TB_CUDA(tb200_launch(swiglu_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, c->stream, gate, up, ld, tokens, k, y, ldy));
    return TB200_OK;
}

int tb200_moe_combine_dev(const float *expert_out, const float *weights, long tokens, int used, long rows, float *out, long ldo) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    if (!expert_out || !weights || !out || tokens < 0 || used < 1 || rows <= 0 || ldo < rows) { tb200_set_error("moe_combine: bad arguments"); return TB200_ERR_INVALID; }
    if (tokens == 0) return TB200_OK;
    const long n = tokens * rows;
    TB_CUDA(tb200_launch(moe_combine_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, c->stream, expert_out, weights, tokens, used, rows, out, ldo));
    return TB200_OK;
}

// norm / swiglu straight into resident activations: the transform lands in the context's scratch arena, the quantizer reads it
int tb200_acts_quantize_rmsnorm(tb200_acts *a, const float *x, long ldx, long tokens, const float *norm_weight, float eps, long k) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    float *tmp = (float *)tb200_scratch2(c, (size_t)tokens * k * 4);
    if (!tmp) return TB200_ERR_CUDA;
    int rc = tb200_rms_norm_mul_dev(x, ldx, norm_weight, eps, tokens, k, tmp, k);
    if (rc) return rc;
    return tb200_acts_quantize(a, tmp, k, tokens);
}

int tb200_acts_quantize_swiglu(tb200_acts *a, const float *gate, const float *up, long ld, long tokens, long k) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    float *tmp = (float *)tb200_scratch2(c, (size_t)tokens * k * 4);
    if (!tmp) return TB200_ERR_CUDA;
    int rc = tb200_silu_mul_dev(gate, up, ld, tokens, k, tmp, k);
    if (rc) return rc;
    return tb200_acts_quantize(a, tmp, k, tokens);
}

}  // extern "C"
