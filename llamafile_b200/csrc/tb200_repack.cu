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
