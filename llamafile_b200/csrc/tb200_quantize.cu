// Activation quantizers, bit-exact to the reference.
//
//   quantize_row_q8_0 : llama.cpp/ggml/src/ggml-quants.c:940-990 (the x86 body llamafile runs):
//       d = amax/127 (stored fp16, RNE); id = amax ? 127/amax : 0 (IEEE divide); q = rne(x*id)
//   quantize_row_q8_K : ggml-quants.c:3593-3631 with llamafile's block layout {d, bsums, qs}
//       (llama.cpp.patches/patches/ggml-common.h.patch:25-35):
//       max = signed value at the FIRST arg-max |x|; iscale = -127/max; q = min(127, nearest_int(iscale*x));
//       bsums[j] = sum of 16 q; d = 1/iscale; all-zero block -> d = 0, q = 0
//   ggml_fp32_to_fp16_row : RNE
//
// All float ops use the explicit _rn intrinsics so that neither -use_fast_math nor FMA contraction
// can change a bit. NaN/Inf inputs are outside the contract (the reference's SIMD max/round
// behaviour on them is ISA-specific).
#include "tb200_common.cuh"

namespace {

// _mm256_cvtps_epi32 + packs (ggml-quants.c:962-977): RNE; NaN / out of int32 range -> INT_MIN; saturate to int8.
// Only differs from a plain round when id = 127/amax overflowed (amax < ~3.7e-37): the reference then stores -128.
__device__ __forceinline__ int cvt_rne_sat8(float v) {
    if (!(v >= -2147483648.f && v < 2147483648.f)) return -128;
    return max(-128, min(127, __float2int_rn(v)));
}

// One 8-lane group per 32-element block: lane g8 holds elements 4*g8 .. 4*g8+3 (one float4).
template <bool kVec>
__global__ void __launch_bounds__(256) quantize_q8_0_kernel(const float *__restrict__ x, long ldx, long rows, long nb,
                                                            uint8_t *__restrict__ raw, int8_t *__restrict__ q,
                                                            int2 *__restrict__ ds) {
    pdl_launch_dependents();
    pdl_wait();   // x comes from a predecessor; the outputs may still be read by one
    const long group = (blockIdx.x * (long)blockDim.x + threadIdx.x) >> 3;
    const int g8 = threadIdx.x & 7;
    const bool live = group < rows * nb;   // whole 8-lane groups go idle together; shuffles stay in-group
    const long row = live ? group / nb : 0, b = live ? group % nb : 0;
    const float *src = x + row * ldx + b * QK + 4 * g8;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (live) {
        if (kVec) v = *reinterpret_cast<const float4 *>(src);
        else v = make_float4(src[0], src[1], src[2], src[3]);
    }
    float amax = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w)));
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    const float d = __fdiv_rn(amax, 127.f);
    const float id = (amax != 0.0f) ? __fdiv_rn(127.f, amax) : 0.0f;
    const int q0 = cvt_rne_sat8(__fmul_rn(v.x, id)), q1 = cvt_rne_sat8(__fmul_rn(v.y, id));
    const int q2 = cvt_rne_sat8(__fmul_rn(v.z, id)), q3 = cvt_rne_sat8(__fmul_rn(v.w, id));
    const uint32_t packed = (uint32_t)(q0 & 0xff) | ((uint32_t)(q1 & 0xff) << 8) | ((uint32_t)(q2 & 0xff) << 16) | ((uint32_t)(q3 & 0xff) << 24);
    const __half dh = __float2half_rn(d);
    int sum = q0 + q1 + q2 + q3;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (!live) return;
    if (raw) {
        uint8_t *blk = raw + (row * nb + b) * sizeof(tb200_block_q8_0);   // 34-byte blocks: 2-byte aligned only
        uint16_t *p = reinterpret_cast<uint16_t *>(blk + 2 + 4 * g8);
        p[0] = (uint16_t)(packed & 0xffff);
        p[1] = (uint16_t)(packed >> 16);
        if (g8 == 0) *reinterpret_cast<uint16_t *>(blk) = __half_as_ushort(dh);
    }
    if (q) {
        *reinterpret_cast<uint32_t *>(q + (row * nb + b) * QK + 4 * g8) = packed;
        if (g8 == 0) ds[row * nb + b] = make_int2(__float_as_int(__half2float(dh)), sum);
    }
}

__device__ __forceinline__ int nearest_int_magic(float fval) {   // ggml-quants.c:1632-1637
    // iscale = -127/max overflows to inf for |max| < ~3.7e-37; the x86 bit pattern of inf/NaN + magic then
    // yields a value whose low byte is 0 (x86 NaN = 0xFFC00000, CUDA NaN = 0x7FFFFFFF would give 127)
    if (!(fabsf(fval) <= 3.4028234664e38f)) return 0;
    const float val = __fadd_rn(fval, 12582912.f);
    return (__float_as_int(val) & 0x007fffff) - 0x00400000;
}

// One warp per 256-element super-block: lane l holds elements 4l..4l+3 and 128+4l..128+4l+3.
template <bool kVec>
__global__ void __launch_bounds__(256) quantize_q8_K_kernel(const float *__restrict__ x, long ldx, long rows, long nb,
                                                            uint8_t *__restrict__ raw, int8_t *__restrict__ q,
                                                            int2 *__restrict__ ds) {
    pdl_launch_dependents();
    pdl_wait();
    const long sb = (blockIdx.x * (long)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (sb >= rows * nb) return;   // warp-uniform
    const long row = sb / nb, b = sb % nb;
    const float *src = x + row * ldx + b * QK_K;
    float v[8];
    if (kVec) {
        const float4 a = *reinterpret_cast<const float4 *>(src + 4 * lane), c = *reinterpret_cast<const float4 *>(src + 128 + 4 * lane);
        v[0] = a.x, v[1] = a.y, v[2] = a.z, v[3] = a.w, v[4] = c.x, v[5] = c.y, v[6] = c.z, v[7] = c.w;
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) v[i] = src[4 * lane + i], v[4 + i] = src[128 + 4 * lane + i];
    }
    // first arg-max |x| (strict >), carrying the signed value
    float amax = 0.f, vmax = 0.f;
    int imax = QK_K;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float ax = fabsf(v[i]);
        const int idx = (i < 4) ? 4 * lane + i : 128 + 4 * lane + (i - 4);
        if (ax > amax) amax = ax, vmax = v[i], imax = idx;   // per-lane indices ascend with i
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float oa = __shfl_xor_sync(0xffffffffu, amax, o), ov = __shfl_xor_sync(0xffffffffu, vmax, o);
        const int oi = __shfl_xor_sync(0xffffffffu, imax, o);
        if (oa > amax || (oa == amax && oi < imax)) amax = oa, vmax = ov, imax = oi;
    }
    int qv[8];
    float d = 0.f;
    if (amax != 0.f) {
        const float iscale = __fdiv_rn(-127.f, vmax);
#pragma unroll
        for (int i = 0; i < 8; ++i) qv[i] = min(127, nearest_int_magic(__fmul_rn(iscale, v[i])));
        d = __fdiv_rn(1.f, iscale);
    } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) qv[i] = 0;
    }
    auto pack = [](const int *p) {
        return (uint32_t)(p[0] & 0xff) | ((uint32_t)(p[1] & 0xff) << 8) | ((uint32_t)(p[2] & 0xff) << 16) | ((uint32_t)(p[3] & 0xff) << 24);
    };
    const uint32_t lo = pack(qv), hi = pack(qv + 4);
    int s_lo = qv[0] + qv[1] + qv[2] + qv[3], s_hi = qv[4] + qv[5] + qv[6] + qv[7];
    s_lo += __shfl_xor_sync(0xffffffffu, s_lo, 1), s_hi += __shfl_xor_sync(0xffffffffu, s_hi, 1);
    s_lo += __shfl_xor_sync(0xffffffffu, s_lo, 2), s_hi += __shfl_xor_sync(0xffffffffu, s_hi, 2);   // 16-element sums
    const int b16_lo = s_lo, b16_hi = s_hi;
    s_lo += __shfl_xor_sync(0xffffffffu, s_lo, 4), s_hi += __shfl_xor_sync(0xffffffffu, s_hi, 4);   // 32-element sums
    if (raw) {
        uint8_t *blk = raw + (row * nb + b) * sizeof(tb200_block_q8_K);   // 292-byte blocks: 4-byte aligned
        if (lane == 0) *reinterpret_cast<float *>(blk) = d;
        int16_t *bs = reinterpret_cast<int16_t *>(blk + 4);
        if ((lane & 3) == 0) bs[lane >> 2] = (int16_t)b16_lo, bs[8 + (lane >> 2)] = (int16_t)b16_hi;
        uint32_t *qs = reinterpret_cast<uint32_t *>(blk + 36);
        qs[lane] = lo, qs[32 + lane] = hi;
    }
    if (q) {
        uint32_t *qq = reinterpret_cast<uint32_t *>(q + (row * nb + b) * QK_K);
        qq[lane] = lo, qq[32 + lane] = hi;
        if ((lane & 7) == 0) {
            int2 *o = ds + (row * nb + b) * 8;
            o[lane >> 3] = make_int2(__float_as_int(d), s_lo);
            o[4 + (lane >> 3)] = make_int2(__float_as_int(d), s_hi);
        }
    }
}

__global__ void f32_to_f16_kernel(const float *__restrict__ x, long ldx, long rows, long k, __half *__restrict__ y) {
    pdl_launch_dependents();
    pdl_wait();
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i >= rows * k) return;
    y[i] = __float2half_rn(x[(i / k) * ldx + (i % k)]);
}

// raw GGUF activation blocks (as a ggml caller hands them to llamafile_sgemm) -> planar
__global__ void unpack_q8_0_kernel(const uint8_t *__restrict__ raw, long ldb, long rows, long nb, int8_t *__restrict__ q,
                                   int2 *__restrict__ ds) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;   // one thread per 4 quants
    const bool live = i < rows * nb * 8;   // 8-lane groups are live or idle as a whole
    const long blk = live ? i >> 3 : 0, row = blk / nb, b = blk % nb;
    const int g8 = i & 7;
    const uint8_t *src = raw + (row * ldb + b) * sizeof(tb200_block_q8_0);
    const uint16_t *p = reinterpret_cast<const uint16_t *>(src + 2 + 4 * g8);
    const uint32_t packed = (uint32_t)p[0] | ((uint32_t)p[1] << 16);
    int sum = __dp4a((int)packed, 0x01010101, 0);
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (!live) return;
    *reinterpret_cast<uint32_t *>(q + blk * QK + 4 * g8) = packed;
    if (g8 == 0) ds[blk] = make_int2(__float_as_int(__half2float(__ushort_as_half(*reinterpret_cast<const uint16_t *>(src)))), sum);
}

// block_q8_1 {fp16 d; fp16 s; int8 qs[32]} (ggml-common.h:193-203): the quants and d are quantize_row_q8_0's; s = fp16(d*sum q) is
// re-formed by the consumer from {d, sum q}
__global__ void unpack_q8_1_kernel(const uint8_t *__restrict__ raw, long ldb, long rows, long nb, int8_t *__restrict__ q, int2 *__restrict__ ds) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;   // one thread per 4 quants
    const bool live = i < rows * nb * 8;
    const long blk = live ? i >> 3 : 0, row = blk / nb, b = blk % nb;
    const int g8 = i & 7;
    const uint8_t *src = raw + (row * ldb + b) * sizeof(tb200_block_q8_1);
    const uint32_t packed = *reinterpret_cast<const uint32_t *>(src + 4 + 4 * g8);   // 36-byte blocks: 4-byte aligned
    int sum = __dp4a((int)packed, 0x01010101, 0);
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (!live) return;
    *reinterpret_cast<uint32_t *>(q + blk * QK + 4 * g8) = packed;
    if (g8 == 0) ds[blk] = make_int2(__float_as_int(__half2float(__ushort_as_half(*reinterpret_cast<const uint16_t *>(src)))), sum);
}

__global__ void unpack_q8_K_kernel(const uint8_t *__restrict__ raw, long ldb, long rows, long nb, int8_t *__restrict__ q,
                                   int2 *__restrict__ ds) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;   // one thread per 4 quants
    if (i >= rows * nb * 64) return;
    const long blk = i >> 6, row = blk / nb, b = blk % nb;
    const int w = i & 63;
    const uint8_t *src = raw + (row * ldb + b) * sizeof(tb200_block_q8_K);
    reinterpret_cast<uint32_t *>(q + blk * QK_K)[w] = reinterpret_cast<const uint32_t *>(src + 36)[w];
    if (w < 8) {
        const int16_t *bs = reinterpret_cast<const int16_t *>(src + 4);
        ds[blk * 8 + w] = make_int2(__float_as_int(*reinterpret_cast<const float *>(src)), (int)bs[2 * w] + (int)bs[2 * w + 1]);
    }
}

bool aligned16(const void *p, long ld_elems) { return (((uintptr_t)p) & 15) == 0 && (ld_elems & 3) == 0; }

}  // namespace

int tb200_launch_quantize_q8_0(tb200_ctx *c, const float *x, long ldx, long rows, long k, void *raw, int8_t *q, int2 *ds) {
    if (k % QK) { tb200_set_error("quantize_q8_0: k=%ld not a multiple of 32", k); return TB200_ERR_INVALID; }
    const long nb = k / QK, threads = rows * nb * 8;
    if (threads == 0) return TB200_OK;
    const unsigned grid = (unsigned)((threads + 255) / 256);
    if (aligned16(x, ldx)) TB_CUDA(tb200_launch(quantize_q8_0_kernel<true>, dim3(grid), dim3(256), 0, c->stream, x, ldx, rows, nb, (uint8_t *)raw, q, ds));
    else TB_CUDA(tb200_launch(quantize_q8_0_kernel<false>, dim3(grid), dim3(256), 0, c->stream, x, ldx, rows, nb, (uint8_t *)raw, q, ds));
    return TB200_OK;
}

int tb200_launch_quantize_q8_K(tb200_ctx *c, const float *x, long ldx, long rows, long k, void *raw, int8_t *q, int2 *ds) {
    if (k % QK_K) { tb200_set_error("quantize_q8_K: k=%ld not a multiple of 256", k); return TB200_ERR_INVALID; }
    const long nb = k / QK_K, threads = rows * nb * 32;
    if (threads == 0) return TB200_OK;
    const unsigned grid = (unsigned)((threads + 255) / 256);
    if (aligned16(x, ldx)) TB_CUDA(tb200_launch(quantize_q8_K_kernel<true>, dim3(grid), dim3(256), 0, c->stream, x, ldx, rows, nb, (uint8_t *)raw, q, ds));
    else TB_CUDA(tb200_launch(quantize_q8_K_kernel<false>, dim3(grid), dim3(256), 0, c->stream, x, ldx, rows, nb, (uint8_t *)raw, q, ds));
    return TB200_OK;
}

int tb200_launch_f32_to_f16(tb200_ctx *c, const float *x, long ldx, long rows, long k, __half *y) {
    const long n = rows * k;
    if (n == 0) return TB200_OK;
    TB_CUDA(tb200_launch(f32_to_f16_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, c->stream, x, ldx, rows, k, y));
    return TB200_OK;
}

int tb200_launch_unpack_q8_0(tb200_ctx *c, const void *raw, long ldb, long rows, long k, int8_t *q, int2 *ds) {
    const long nb = k / QK, n = rows * nb * 8;
    if (n == 0) return TB200_OK;
    unpack_q8_0_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>((const uint8_t *)raw, ldb, rows, nb, q, ds);
    tb200_count_launch();
    TB_CUDA(cudaGetLastError());
    return TB200_OK;
}

int tb200_launch_unpack_q8_K(tb200_ctx *c, const void *raw, long ldb, long rows, long k, int8_t *q, int2 *ds) {
    const long nb = k / QK_K, n = rows * nb * 64;
    if (n == 0) return TB200_OK;
    unpack_q8_K_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>((const uint8_t *)raw, ldb, rows, nb, q, ds);
    tb200_count_launch();
    TB_CUDA(cudaGetLastError());
    return TB200_OK;
}

int tb200_launch_unpack_q8_1(tb200_ctx *c, const void *raw, long ldb, long rows, long k, int8_t *q, int2 *ds) {
    const long nb = k / QK, n = rows * nb * 8;
    if (n == 0) return TB200_OK;
    unpack_q8_1_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>((const uint8_t *)raw, ldb, rows, nb, q, ds);
    tb200_count_launch();
    TB_CUDA(cudaGetLastError());
    return TB200_OK;
}
