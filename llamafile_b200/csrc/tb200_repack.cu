// One-time upload repack: GGUF rows (block structs, 2-byte aligned) -> 16-byte-aligned planes.
// The GGUF bytes are the wire/storage format and stay unchanged; this only separates each block's
// scale header from its quants so decode can issue 128-bit coalesced loads and the prefill path can
// address the quants with TMA (global strides must be multiples of 16 bytes). Same total bytes.
//   block_q4_0 {fp16 d; u8 qs[16]}            (ggml-common.h:144-149) -> sc[b] = d, qs[b] = 16 B
//   block_q8_0 {fp16 d; i8 qs[32]}            (ggml-common.h:186-191) -> sc[b] = d, qs[b] = 32 B
//   block_q4_K {d, dmin, scales[12]; qs[128]} (ggml-common.h:266-277) -> sc[b] = first 16 B, qs[b] = 128 B
//   block_q6_K {ql[128]; qh[64]; scales[16]; d} (ggml-common.h:299-307) -> qs[b] = 4 units x 48 B, sc[b] = 16 B, d[b]
#include "tb200_common.cuh"

namespace {

__global__ void repack_q4_0_kernel(const uint8_t *__restrict__ raw, size_t row_stride, long rows, long nb,
                                   uint4 *__restrict__ qs, uint16_t *__restrict__ sc) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i >= rows * nb) return;
    const long row = i / nb, b = i % nb;
    const uint16_t *p = reinterpret_cast<const uint16_t *>(raw + row * row_stride + b * sizeof(tb200_block_q4_0));
    uint16_t h[9];
#pragma unroll
    for (int j = 0; j < 9; ++j) h[j] = p[j];
    sc[i] = h[0];
    qs[i] = make_uint4(h[1] | ((uint32_t)h[2] << 16), h[3] | ((uint32_t)h[4] << 16), h[5] | ((uint32_t)h[6] << 16),
                       h[7] | ((uint32_t)h[8] << 16));
}

__global__ void repack_q8_0_kernel(const uint8_t *__restrict__ raw, size_t row_stride, long rows, long nb,
                                   uint4 *__restrict__ qs, uint16_t *__restrict__ sc) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i >= rows * nb) return;
    const long row = i / nb, b = i % nb;
    const uint16_t *p = reinterpret_cast<const uint16_t *>(raw + row * row_stride + b * sizeof(tb200_block_q8_0));
    uint16_t h[17];
#pragma unroll
    for (int j = 0; j < 17; ++j) h[j] = p[j];
    sc[i] = h[0];
    qs[2 * i] = make_uint4(h[1] | ((uint32_t)h[2] << 16), h[3] | ((uint32_t)h[4] << 16), h[5] | ((uint32_t)h[6] << 16),
                           h[7] | ((uint32_t)h[8] << 16));
    qs[2 * i + 1] = make_uint4(h[9] | ((uint32_t)h[10] << 16), h[11] | ((uint32_t)h[12] << 16),
                               h[13] | ((uint32_t)h[14] << 16), h[15] | ((uint32_t)h[16] << 16));
}

// one thread per 32-bit word of a 144-byte block (36 words)
__global__ void repack_q4_K_kernel(const uint8_t *__restrict__ raw, size_t row_stride, long rows, long nb,
                                   uint32_t *__restrict__ qs, uint32_t *__restrict__ sc) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i >= rows * nb * 36) return;
    const long blk = i / 36;
    const int w = (int)(i % 36);
    const long row = blk / nb, b = blk % nb;
    const uint8_t *src = raw + row * row_stride + b * sizeof(tb200_block_q4_K) + 4 * w;
    uint32_t v;
    if ((((uintptr_t)src) & 3) == 0) v = *reinterpret_cast<const uint32_t *>(src);
    else v = (uint32_t)src[0] | ((uint32_t)src[1] << 8) | ((uint32_t)src[2] << 16) | ((uint32_t)src[3] << 24);
    if (w < 4) sc[blk * 4 + w] = v;
    else qs[blk * 32 + (w - 4)] = v;
}

// block_q6_K {ql[128], qh[64], scales[16], d} (ggml-common.h:299-307) -> 4 units of 48 B + scales + d. One thread per
// 16-byte piece of the 192 quant bytes (12 pieces) plus one for the 16 scales and d. Unit u = 2*half + r covers
// l in [16r, 16r+16) of that half, i.e. the four 16-element groups j = 0..3 at elements 128*half + 32*j + 16*r
// (the values dequantize_row_q6_K, ggml-quants.c:2970-3000, calls q1..q4). The 6-bit values are re-cut so that a
// decode lane needs no nibble/high-bit merge: piece j (j = 0,1,2), byte e = q_j[e] << 2 | bits (5-2j, 4-2j) of q_3[e].
// (piece & 0xfc) ^ 0x80 is then the int8 4*(q_j - 32), and the three 2-bit tails reassemble 4*q_3.
__global__ void repack_q6_K_kernel(const uint8_t *__restrict__ raw, size_t row_stride, long rows, long nb,
                                   uint8_t *__restrict__ qs, uint8_t *__restrict__ sc, uint16_t *__restrict__ d) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i >= rows * nb * 13) return;
    const long blk = i / 13;
    const int piece = (int)(i % 13);
    const long row = blk / nb, b = blk % nb;
    const uint8_t *src = raw + row * row_stride + b * sizeof(tb200_block_q6_K);
    if (piece == 12) {   // scales regrouped per unit: byte 4u + j = scales[8*half + r + 2j]
        for (int u = 0; u < 4; ++u)
            for (int j = 0; j < 4; ++j) sc[blk * 16 + 4 * u + j] = src[192 + 8 * (u >> 1) + (u & 1) + 2 * j];
        d[blk] = (uint16_t)src[208] | ((uint16_t)src[209] << 8);
        return;
    }
    const int u = piece / 3, part = piece % 3, half = u >> 1, r = u & 1;
    const uint8_t *ql = src + 64 * half + 16 * r, *qh = src + 128 + 32 * half + 16 * r;
    uint8_t *dst = qs + blk * 192 + piece * 16;
    for (int e = 0; e < 16; ++e) {
        const int h = qh[e];
        const int q3 = (ql[e + 32] >> 4) | (((h >> 6) & 3) << 4);
        const int qj = part == 0 ? ((ql[e] & 0xF) | ((h & 3) << 4)) : part == 1 ? ((ql[e + 32] & 0xF) | (((h >> 2) & 3) << 4)) : ((ql[e] >> 4) | (((h >> 4) & 3) << 4));
        dst[e] = (uint8_t)((qj << 2) | ((q3 >> (4 - 2 * part)) & 3));
    }
}

// ---- the generic form G8 (tb200_common.cuh): every 32-block as int8 quants + fp16 {a, b}, value = a*q + b. One thread per 32-block.
//   block_q4_1  (ggml-common.h:151-161, dequantize_row_q4_1 ggml-quants.c:1535-1554): x = d*q + m
//   block_q5_0  (ggml-common.h:164-169, dequantize_row_q5_0 :1556-1580):              x = d*(q5 - 16), 5th bits in qh
//   block_q5_K  (ggml-common.h:280-295, dequantize_row_q5_K :2756-2781):              x = (d*sc)*q5 - dmin*m per 32-sub-block
//   block_iq4_xs (ggml-common.h:393-398, dequantize_row_iq4_xs :3568-3590):           x = d*(ls - 32) * kvalues_iq4nl[q4]
__device__ __constant__ int8_t kvalues_iq4nl_dev[16] = {-127, -104, -83, -65, -49, -35, -22, -10, 1, 13, 25, 38, 53, 69, 89, 113};   // ggml-quants.c:3548

__device__ __forceinline__ void g8_store(int8_t *qs, uint32_t *sc, long blk, const int (&q)[32], float a, float b) {
    uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
        w[i] = (uint32_t)(q[4 * i] & 0xff) | ((uint32_t)(q[4 * i + 1] & 0xff) << 8) | ((uint32_t)(q[4 * i + 2] & 0xff) << 16) | ((uint32_t)(q[4 * i + 3] & 0xff) << 24);
    uint4 *dst = reinterpret_cast<uint4 *>(qs + blk * 32);
    dst[0] = make_uint4(w[0], w[1], w[2], w[3]);
    dst[1] = make_uint4(w[4], w[5], w[6], w[7]);
    sc[blk] = (uint32_t)__half_as_ushort(__float2half_rn(a)) | ((uint32_t)__half_as_ushort(__float2half_rn(b)) << 16);
}
__device__ __forceinline__ float h16(const uint8_t *p) { return __half2float(__ushort_as_half((unsigned short)(p[0] | (p[1] << 8)))); }

__global__ void repack_g8_kernel(int type, const uint8_t *__restrict__ raw, size_t row_stride, long rows, long nb32, int8_t *__restrict__ qs, uint32_t *__restrict__ sc) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;   // 32-block index
    if (i >= rows * nb32) return;
    const long row = i / nb32, b = i % nb32;
    const uint8_t *r = raw + row * row_stride;
    int q[32];
    float a, bb = 0.f;
    if (type == TB200_TYPE_Q4_1) {
        const uint8_t *x = r + b * sizeof(tb200_block_q4_1);
        a = h16(x), bb = h16(x + 2);
        for (int j = 0; j < 16; ++j) q[j] = x[4 + j] & 0xF, q[j + 16] = x[4 + j] >> 4;
    } else if (type == TB200_TYPE_Q5_0) {
        const uint8_t *x = r + b * sizeof(tb200_block_q5_0);
        a = h16(x);
        const uint32_t qh = (uint32_t)x[2] | ((uint32_t)x[3] << 8) | ((uint32_t)x[4] << 16) | ((uint32_t)x[5] << 24);
        for (int j = 0; j < 16; ++j) {
            q[j] = (int)((x[6 + j] & 0x0F) | (((qh >> j) << 4) & 0x10)) - 16;
            q[j + 16] = (int)((x[6 + j] >> 4) | ((qh >> (j + 12)) & 0x10)) - 16;
        }
    } else if (type == TB200_TYPE_Q5_K) {
        const long sb = b >> 3;
        const int s = (int)(b & 7);
        const uint8_t *x = r + sb * sizeof(tb200_block_q5_K);
        const uint8_t *scales = x + 4, *qh = x + 16, *ql = x + 48 + 32 * (s >> 1);
        int scv, mv;   // get_scale_min_k4, ggml-quants.c:1891-1899
        if (s < 4) scv = scales[s] & 63, mv = scales[s + 4] & 63;
        else scv = (scales[s + 4] & 0xF) | ((scales[s - 4] >> 6) << 4), mv = (scales[s + 4] >> 4) | ((scales[s] >> 6) << 4);
        a = h16(x) * (float)scv, bb = -(h16(x + 2) * (float)mv);
        for (int l = 0; l < 32; ++l) q[l] = (int)((s & 1) ? (ql[l] >> 4) : (ql[l] & 0xF)) + (((qh[l] >> s) & 1) ? 16 : 0);
    } else {   // IQ4_XS
        const long sb = b >> 3;
        const int ib = (int)(b & 7);
        const uint8_t *x = r + sb * sizeof(tb200_block_iq4_xs);
        const uint32_t scales_h = (uint32_t)x[2] | ((uint32_t)x[3] << 8);
        const int ls = ((x[4 + ib / 2] >> (4 * (ib % 2))) & 0xf) | (((scales_h >> (2 * ib)) & 3) << 4);
        a = h16(x) * (float)(ls - 32);
        const uint8_t *qsrc = x + 8 + 16 * ib;
        for (int j = 0; j < 16; ++j) q[j] = kvalues_iq4nl_dev[qsrc[j] & 0xf], q[j + 16] = kvalues_iq4nl_dev[qsrc[j] >> 4];
    }
    g8_store(qs, sc, i, q, a, bb);
}

}  // namespace

int tb200_launch_repack(tb200_ctx *c, const void *raw, int type, long rows, long k, size_t row_stride, void *qs, void *sc, void *d) {
    if (rows == 0 || k == 0) return TB200_OK;
    const uint8_t *r = (const uint8_t *)raw;
    switch (type) {
    case TB200_TYPE_Q4_0: {
        const long nb = k / QK, n = rows * nb;
        repack_q4_0_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(r, row_stride, rows, nb, (uint4 *)qs, (uint16_t *)sc);
        break;
    }
    case TB200_TYPE_Q8_0: {
        const long nb = k / QK, n = rows * nb;
        repack_q8_0_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(r, row_stride, rows, nb, (uint4 *)qs, (uint16_t *)sc);
        break;
    }
    case TB200_TYPE_Q4_K: {
        const long nb = k / QK_K, n = rows * nb * 36;
        repack_q4_K_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(r, row_stride, rows, nb, (uint32_t *)qs, (uint32_t *)sc);
        break;
    }
    case TB200_TYPE_Q6_K: {
        const long nb = k / QK_K, n = rows * nb * 13;
        repack_q6_K_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(r, row_stride, rows, nb, (uint8_t *)qs, (uint8_t *)sc, (uint16_t *)d);
        break;
    }
    case TB200_TYPE_Q4_1: case TB200_TYPE_Q5_0: case TB200_TYPE_Q5_K: case TB200_TYPE_IQ4_XS: {
        const long nb32 = k / QK, n = rows * nb32;
        repack_g8_kernel<<<(unsigned)((n + 127) / 128), 128, 0, c->stream>>>(type, r, row_stride, rows, nb32, (int8_t *)qs, (uint32_t *)sc);
        break;
    }
    case TB200_TYPE_F16:
        TB_CUDA(cudaMemcpy2DAsync(qs, (size_t)k * 2, raw, row_stride, (size_t)k * 2, (size_t)rows, cudaMemcpyDeviceToDevice, c->stream));
        return TB200_OK;
    default:
        tb200_set_error("repack: unsupported weight type %d", type);
        return TB200_ERR_UNSUPPORTED;
    }
    tb200_count_launch();
    TB_CUDA(cudaGetLastError());
    return TB200_OK;
}
