/* TEST INFRASTRUCTURE — see tb_oracle.h. Plain scalar C; every function cites the reference
 * lines it restates (paths relative to /root/reference). Compiled with -ffp-contract=off so the
 * float arithmetic below has source-level IEEE semantics. */
#include "tb_oracle.h"

#include <math.h>
#include <string.h>

#define QK 32
#define QK_K 256

/* llama.cpp/ggml/src/ggml-common.h:144-149, 186-191, 266-277; block_q8_K :311-315 as patched */
#pragma pack(push, 1)
typedef struct { uint16_t d; uint8_t qs[16]; } blk_q4_0;                       /* 18 B */
typedef struct { uint16_t d; int8_t qs[32]; } blk_q8_0;                        /* 34 B */
typedef struct { uint16_t d, dmin; uint8_t scales[12]; uint8_t qs[128]; } blk_q4_K; /* 144 B */
/* llamafile swaps bsums and qs relative to upstream (llama.cpp.patches/patches/ggml-common.h.patch:25-35) */
typedef struct { float d; int16_t bsums[16]; int8_t qs[256]; } blk_q8_K;       /* 292 B */
/* llama.cpp/ggml/src/ggml-common.h:299-307 */
typedef struct { uint8_t ql[128]; uint8_t qh[64]; int8_t scales[16]; uint16_t d; } blk_q6_K; /* 210 B */
/* the remaining types iqk serves on the same entry points (ggml-common.h:151-169, 193-203, 280-295, 393-398) */
typedef struct { uint16_t d, m; uint8_t qs[16]; } blk_q4_1;                                   /* 20 B */
typedef struct { uint16_t d; uint8_t qh[4]; uint8_t qs[16]; } blk_q5_0;                        /* 22 B */
typedef struct { uint16_t d, s; int8_t qs[32]; } blk_q8_1;                                     /* 36 B */
typedef struct { uint16_t d, dmin; uint8_t scales[12]; uint8_t qh[32]; uint8_t qs[128]; } blk_q5_K;   /* 176 B */
typedef struct { uint16_t d; uint16_t scales_h; uint8_t scales_l[4]; uint8_t qs[128]; } blk_iq4_xs;   /* 136 B */
static const int8_t kvalues_iq4nl[16] = {-127, -104, -83, -65, -49, -35, -22, -10, 1, 13, 25, 38, 53, 69, 89, 113};   /* ggml-quants.c:3548 */
#pragma pack(pop)
_Static_assert(sizeof(blk_q6_K) == 210, "q6_K");
_Static_assert(sizeof(blk_q4_0) == 18, "q4_0");
_Static_assert(sizeof(blk_q8_0) == 34, "q8_0");
_Static_assert(sizeof(blk_q4_K) == 144, "q4_K");
_Static_assert(sizeof(blk_q8_K) == 292, "q8_K");

/* IEEE binary16 <-> binary32, round-to-nearest-even (what F16C / ggml_compute_fp32_to_fp16 do,
 * ggml-impl.h:491-497). */
static inline float h2f(uint16_t h) { _Float16 v; memcpy(&v, &h, 2); return (float)v; }
static inline uint16_t f2h(float f) { _Float16 v = (_Float16)f; uint16_t h; memcpy(&h, &v, 2); return h; }

size_t oracle_type_size(int type) {
    switch (type) {
    case OR_F32: return 4; case OR_F16: return 2; case OR_Q4_0: return 18; case OR_Q8_0: return 34;
    case OR_Q4_K: return 144; case OR_Q6_K: return 210; case OR_Q8_K: return 292; case OR_I32: return 4;
    case OR_Q4_1: return 20; case OR_Q5_0: return 22; case OR_Q8_1: return 36; case OR_Q5_K: return 176; case OR_IQ4_XS: return 136; default: return 0;
    }
}
int oracle_blck_size(int type) {
    switch (type) {
    case OR_F32: case OR_F16: case OR_I32: return 1; case OR_Q4_0: case OR_Q8_0: case OR_Q4_1: case OR_Q5_0: case OR_Q8_1: return 32;
    case OR_Q4_K: case OR_Q6_K: case OR_Q8_K: case OR_Q5_K: case OR_IQ4_XS: return 256; default: return 0;
    }
}

/* ---------------------------------------------------------------- quantizers */

/* _mm256_round_ps(NEAREST) + _mm256_cvtps_epi32 + packs_epi32 + packs_epi16 (ggml-quants.c:962-977):
 * RNE, NaN or out-of-int32-range -> 0x80000000 ("integer indefinite"), then signed saturation to int8. */
static inline int8_t cvt_rne_sat8(float v) {
    if (!(v >= -2147483648.f && v < 2147483648.f)) return -128;
    long r = lrintf(v);
    return (int8_t)(r < -128 ? -128 : r > 127 ? 127 : r);
}

/* ggml-quants.c:940-990 (the x86 AVX/AVX2 body that runs at llamafile run time):
 * d = amax/127 stored as fp16; id = 127/amax (IEEE divide; overflows to inf for amax < ~3.7e-37);
 * q = sat8(rne(x*id)). */
void oracle_quantize_row_q8_0(const float *x, void *vy, int64_t k) {
    blk_q8_0 *y = (blk_q8_0 *)vy;
    for (int64_t i = 0; i < k / QK; ++i, x += QK) {
        float amax = 0.0f;
        for (int j = 0; j < QK; ++j) { float a = fabsf(x[j]); if (a > amax) amax = a; }
        const float d = amax / 127.f;
        const float id = (amax != 0.0f) ? 127.f / amax : 0.0f;
        y[i].d = f2h(d);
        for (int j = 0; j < QK; ++j) y[i].qs[j] = cvt_rne_sat8(x[j] * id);
    }
}

/* quantize_row_q8_1, x86 runtime body (ggml-quants.c:1276-1321): quantize_row_q8_0's d and quants, plus s = fp16(d * sum q)
 * with the UNROUNDED d. */
void oracle_quantize_row_q8_1(const float *x, void *vy, int64_t k) {
    blk_q8_1 *y = (blk_q8_1 *)vy;
    for (int64_t i = 0; i < k / QK; ++i, x += QK) {
        float amax = 0.0f;
        for (int j = 0; j < QK; ++j) { float a = fabsf(x[j]); if (a > amax) amax = a; }
        const float d = amax / 127.f;
        const float id = (amax != 0.0f) ? 127.f / amax : 0.0f;
        y[i].d = f2h(d);
        int sum = 0;
        for (int j = 0; j < QK; ++j) { y[i].qs[j] = cvt_rne_sat8(x[j] * id); sum += y[i].qs[j]; }
        y[i].s = f2h(d * sum);
    }
}

/* ggml-quants.c:841-864: d = amax/127; id = 1/d; q = roundf(x*id) (half away from zero). */
void oracle_quantize_row_q8_0_ref(const float *x, void *vy, int64_t k) {
    blk_q8_0 *y = (blk_q8_0 *)vy;
    for (int64_t i = 0; i < k / QK; ++i, x += QK) {
        float amax = 0.0f;
        for (int j = 0; j < QK; ++j) { float a = fabsf(x[j]); if (a > amax) amax = a; }
        const float d = amax / 127;
        const float id = d ? 1.0f / d : 0.0f;
        y[i].d = f2h(d);
        for (int j = 0; j < QK; ++j) y[i].qs[j] = (int8_t)roundf(x[j] * id);
    }
}

/* ggml-quants.c:1632-1637 */
static inline int nearest_int(float fval) {
    float val = fval + 12582912.f;
    int i; memcpy(&i, &val, sizeof(int));
    return (i & 0x007fffff) - 0x00400000;
}

/* ggml-quants.c:3593-3631 (quantize_row_q8_K == _ref, :3643-3645). An all-zero block writes d=0 and
 * qs=0; the reference leaves bsums untouched there, we write 0 (tests pre-zero the destination). */
void oracle_quantize_row_q8_K(const float *x, void *vy, int64_t k) {
    blk_q8_K *y = (blk_q8_K *)vy;
    for (int64_t i = 0; i < k / QK_K; ++i, x += QK_K) {
        float max = 0, amax = 0;
        for (int j = 0; j < QK_K; ++j) { float ax = fabsf(x[j]); if (ax > amax) { amax = ax; max = x[j]; } }
        if (!amax) { memset(&y[i], 0, sizeof(y[i])); continue; }
        const float iscale = -127.f / max;
        for (int j = 0; j < QK_K; ++j) { int v = nearest_int(iscale * x[j]); y[i].qs[j] = (int8_t)(v < 127 ? v : 127); }
        for (int j = 0; j < QK_K / 16; ++j) {
            int sum = 0;
            for (int ii = 0; ii < 16; ++ii) sum += y[i].qs[j * 16 + ii];
            y[i].bsums[j] = (int16_t)sum;
        }
        y[i].d = 1 / iscale;
    }
}

/* ggml-quants.c:661-695 */
void oracle_quantize_row_q4_0_ref(const float *x, void *vy, int64_t k) {
    blk_q4_0 *y = (blk_q4_0 *)vy;
    for (int64_t i = 0; i < k / QK; ++i, x += QK) {
        float amax = 0.0f, max = 0.0f;
        for (int j = 0; j < QK; ++j) { float v = x[j]; if (amax < fabsf(v)) { amax = fabsf(v); max = v; } }
        const float d = max / -8;
        const float id = d ? 1.0f / d : 0.0f;
        y[i].d = f2h(d);
        for (int j = 0; j < QK / 2; ++j) {
            const float x0 = x[j] * id, x1 = x[QK / 2 + j] * id;
            int a = (int8_t)(x0 + 8.5f), b = (int8_t)(x1 + 8.5f);
            uint8_t xi0 = (uint8_t)(a < 15 ? a : 15), xi1 = (uint8_t)(b < 15 ? b : 15);
            y[i].qs[j] = xi0 | (xi1 << 4);
        }
    }
}

/* ggml.c ggml_fp32_to_fp16_row (F16C: _mm_cvtps_ph RNE) */
void oracle_fp32_to_fp16_row(const float *x, uint16_t *y, int64_t n) { for (int64_t i = 0; i < n; ++i) y[i] = f2h(x[i]); }
void oracle_fp16_to_fp32_row(const uint16_t *x, float *y, int64_t n) { for (int64_t i = 0; i < n; ++i) y[i] = h2f(x[i]); }

/* ggml-quants.c:1891-1899 */
static inline void get_scale_min_k4(int j, const uint8_t *q, uint8_t *d, uint8_t *m) {
    if (j < 4) { *d = q[j] & 63; *m = q[j + 4] & 63; }
    else { *d = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4); *m = (q[j + 4] >> 4) | ((q[j] >> 6) << 4); }
}

/* ggml-quants.c:1515-1533 (q4_0), :1640-ish (q8_0), :2548-2570 (q4_K) */
void oracle_dequantize_row(int type, const void *vx, float *y, int64_t k) {
    if (type == OR_Q4_0) {
        const blk_q4_0 *x = (const blk_q4_0 *)vx;
        for (int64_t i = 0; i < k / QK; ++i) {
            const float d = h2f(x[i].d);
            for (int j = 0; j < 16; ++j) {
                y[i * QK + j] = ((x[i].qs[j] & 0x0F) - 8) * d;
                y[i * QK + j + 16] = ((x[i].qs[j] >> 4) - 8) * d;
            }
        }
    } else if (type == OR_Q8_0) {
        const blk_q8_0 *x = (const blk_q8_0 *)vx;
        for (int64_t i = 0; i < k / QK; ++i) { const float d = h2f(x[i].d); for (int j = 0; j < QK; ++j) y[i * QK + j] = x[i].qs[j] * d; }
    } else if (type == OR_Q4_K) {
        const blk_q4_K *x = (const blk_q4_K *)vx;
        for (int64_t i = 0; i < k / QK_K; ++i) {
            const uint8_t *q = x[i].qs;
            const float d = h2f(x[i].d), min = h2f(x[i].dmin);
            int is = 0; uint8_t sc, m;
            for (int j = 0; j < QK_K; j += 64) {
                get_scale_min_k4(is + 0, x[i].scales, &sc, &m); const float d1 = d * sc, m1 = min * m;
                get_scale_min_k4(is + 1, x[i].scales, &sc, &m); const float d2 = d * sc, m2 = min * m;
                for (int l = 0; l < 32; ++l) *y++ = d1 * (q[l] & 0xF) - m1;
                for (int l = 0; l < 32; ++l) *y++ = d2 * (q[l] >> 4) - m2;
                q += 32; is += 2;
            }
        }
    } else if (type == OR_Q6_K) {   /* ggml-quants.c:2970-3000 */
        const blk_q6_K *x = (const blk_q6_K *)vx;
        for (int64_t i = 0; i < k / QK_K; ++i) {
            const float d = h2f(x[i].d);
            const uint8_t *ql = x[i].ql, *qh = x[i].qh;
            const int8_t *sc = x[i].scales;
            for (int n = 0; n < QK_K; n += 128) {
                for (int l = 0; l < 32; ++l) {
                    const int is = l / 16;
                    const int8_t q1 = (int8_t)((ql[l + 0] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
                    const int8_t q2 = (int8_t)((ql[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
                    const int8_t q3 = (int8_t)((ql[l + 0] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
                    const int8_t q4 = (int8_t)((ql[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
                    y[l + 0] = d * sc[is + 0] * q1;
                    y[l + 32] = d * sc[is + 2] * q2;
                    y[l + 64] = d * sc[is + 4] * q3;
                    y[l + 96] = d * sc[is + 6] * q4;
                }
                y += 128; ql += 64; qh += 32; sc += 8;
            }
        }
    } else if (type == OR_Q4_1) {   /* ggml-quants.c:1535-1554 */
        const blk_q4_1 *x = (const blk_q4_1 *)vx;
        for (int64_t i = 0; i < k / QK; ++i) {
            const float d = h2f(x[i].d), m = h2f(x[i].m);
            for (int j = 0; j < 16; ++j) { y[i * QK + j] = (x[i].qs[j] & 0x0F) * d + m; y[i * QK + j + 16] = (x[i].qs[j] >> 4) * d + m; }
        }
    } else if (type == OR_Q5_0) {   /* ggml-quants.c:1556-1580 */
        const blk_q5_0 *x = (const blk_q5_0 *)vx;
        for (int64_t i = 0; i < k / QK; ++i) {
            const float d = h2f(x[i].d);
            uint32_t qh; memcpy(&qh, x[i].qh, 4);
            for (int j = 0; j < 16; ++j) {
                const uint8_t xh_0 = ((qh >> (j + 0)) << 4) & 0x10, xh_1 = ((qh >> (j + 12))) & 0x10;
                y[i * QK + j] = (((x[i].qs[j] & 0x0F) | xh_0) - 16) * d;
                y[i * QK + j + 16] = (((x[i].qs[j] >> 4) | xh_1) - 16) * d;
            }
        }
    } else if (type == OR_Q5_K) {   /* ggml-quants.c:2756-2781 */
        const blk_q5_K *x = (const blk_q5_K *)vx;
        for (int64_t i = 0; i < k / QK_K; ++i) {
            const uint8_t *ql = x[i].qs, *qh = x[i].qh;
            const float d = h2f(x[i].d), min = h2f(x[i].dmin);
            int is = 0; uint8_t sc, m, u1 = 1, u2 = 2;
            for (int j = 0; j < QK_K; j += 64) {
                get_scale_min_k4(is + 0, x[i].scales, &sc, &m); const float d1 = d * sc, m1 = min * m;
                get_scale_min_k4(is + 1, x[i].scales, &sc, &m); const float d2 = d * sc, m2 = min * m;
                for (int l = 0; l < 32; ++l) *y++ = d1 * ((ql[l] & 0xF) + (qh[l] & u1 ? 16 : 0)) - m1;
                for (int l = 0; l < 32; ++l) *y++ = d2 * ((ql[l] >> 4) + (qh[l] & u2 ? 16 : 0)) - m2;
                ql += 32; is += 2; u1 <<= 2; u2 <<= 2;
            }
        }
    } else if (type == OR_IQ4_XS) {   /* ggml-quants.c:3568-3590 */
        const blk_iq4_xs *x = (const blk_iq4_xs *)vx;
        for (int64_t i = 0; i < k / QK_K; ++i) {
            const uint8_t *qs = x[i].qs;
            const float d = h2f(x[i].d);
            for (int ib = 0; ib < QK_K / 32; ++ib) {
                const int ls = ((x[i].scales_l[ib / 2] >> 4 * (ib % 2)) & 0xf) | (((x[i].scales_h >> 2 * ib) & 3) << 4);
                const float dl = d * (ls - 32);
                for (int j = 0; j < 16; ++j) { y[j] = dl * kvalues_iq4nl[qs[j] & 0xf]; y[j + 16] = dl * kvalues_iq4nl[qs[j] >> 4]; }
                y += 32; qs += 16;
            }
        }
    } else if (type == OR_F16) {
        oracle_fp16_to_fp32_row((const uint16_t *)vx, y, k);
    } else if (type == OR_F32) {
        memcpy(y, vx, (size_t)k * 4);
    }
}

/* ---------------------------------------------------------------- exact integer block dots */

/* ggml-quants.c:4192-4206 (scalar tail of ggml_vec_dot_q4_0_q8_0) */
void oracle_blockdot_q4_0_q8_0(const void *vx, const void *vy, int64_t nb, int32_t *out) {
    const blk_q4_0 *x = (const blk_q4_0 *)vx; const blk_q8_0 *y = (const blk_q8_0 *)vy;
    for (int64_t ib = 0; ib < nb; ++ib) {
        int sumi0 = 0, sumi1 = 0;
        for (int j = 0; j < QK / 2; ++j) {
            const int v0 = (x[ib].qs[j] & 0x0F) - 8, v1 = (x[ib].qs[j] >> 4) - 8;
            sumi0 += v0 * y[ib].qs[j];
            sumi1 += v1 * y[ib].qs[j + QK / 2];
        }
        out[ib] = sumi0 + sumi1;
    }
}

/* ggml-quants.c:5227 ff. scalar tail of ggml_vec_dot_q8_0_q8_0 */
void oracle_blockdot_q8_0_q8_0(const void *vx, const void *vy, int64_t nb, int32_t *out) {
    const blk_q8_0 *x = (const blk_q8_0 *)vx; const blk_q8_0 *y = (const blk_q8_0 *)vy;
    for (int64_t ib = 0; ib < nb; ++ib) {
        int sumi = 0;
        for (int j = 0; j < QK; ++j) sumi += x[ib].qs[j] * y[ib].qs[j];
        out[ib] = sumi;
    }
}

/* ggml-quants.c:7341-7397 (scalar branch of ggml_vec_dot_q4_K_q8_K), integer parts only */
void oracle_blockdot_q4_K_q8_K(const void *vx, const void *vy, int64_t nb, int32_t *out) {
    const blk_q4_K *x = (const blk_q4_K *)vx; const blk_q8_K *y = (const blk_q8_K *)vy;
    for (int64_t i = 0; i < nb; ++i) {
        int32_t sum_q = 0, sum_m = 0;
        for (int s = 0; s < 8; ++s) {
            uint8_t sc, m; get_scale_min_k4(s, x[i].scales, &sc, &m);
            const uint8_t *q4 = x[i].qs + (s / 2) * 32;
            const int8_t *q8 = y[i].qs + s * 32;
            int32_t dot = 0;
            for (int l = 0; l < 32; ++l) dot += (int)((s & 1) ? (q4[l] >> 4) : (q4[l] & 0xF)) * q8[l];
            sum_q += (int32_t)sc * dot;
            sum_m += (int32_t)m * (y[i].bsums[2 * s] + y[i].bsums[2 * s + 1]);
        }
        out[2 * i] = sum_q; out[2 * i + 1] = sum_m;
    }
}

/* ggml-quants.c:8642-8682 (scalar branch of ggml_vec_dot_q6_K_q8_K), integer part: one int32 per super-block,
 * sum over the 16 sub-blocks of 16 of scale_j * sum (q6 - 32) * q8 */
void oracle_blockdot_q6_K_q8_K(const void *vx, const void *vy, int64_t nb, int32_t *out) {
    const blk_q6_K *x = (const blk_q6_K *)vx; const blk_q8_K *y = (const blk_q8_K *)vy;
    for (int64_t i = 0; i < nb; ++i) {
        int8_t a[QK_K];
        const uint8_t *q4 = x[i].ql, *qh = x[i].qh;
        int8_t *ap = a;
        for (int j = 0; j < QK_K; j += 128) {
            for (int l = 0; l < 32; ++l) {
                ap[l + 0] = (int8_t)((q4[l + 0] & 0xF) | (((qh[l] >> 0) & 3) << 4)) - 32;
                ap[l + 32] = (int8_t)((q4[l + 32] & 0xF) | (((qh[l] >> 2) & 3) << 4)) - 32;
                ap[l + 64] = (int8_t)((q4[l + 0] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32;
                ap[l + 96] = (int8_t)((q4[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32;
            }
            ap += 128; q4 += 64; qh += 32;
        }
        int32_t sum = 0;
        for (int j = 0; j < QK_K / 16; ++j) {
            int32_t dot = 0;
            for (int l = 0; l < 16; ++l) dot += (int32_t)a[16 * j + l] * y[i].qs[16 * j + l];
            sum += (int32_t)x[i].scales[j] * dot;
        }
        out[i] = sum;
    }
}

/* ggml_vec_dot_{q4_0_q8_0,q8_0_q8_0,q4_K_q8_K} scalar semantics: per block sumi*d_x*d_y summed in
 * fp32 in block order (ggml-quants.c:4205, :7385-7392). F16: fp32 products, double accumulation
 * (tinyBLAS's ruler-function summation, tinyblas_cpu.h:534-584, keeps fp32 error logarithmic; a
 * double accumulator is the order-free statement of that intent). */
float oracle_vec_dot(int xtype, int64_t n, const void *vx, const void *vy) {
    if (xtype == OR_Q4_0 || xtype == OR_Q8_0) {
        const int64_t nb = n / QK; float sumf = 0;
        const blk_q8_0 *y = (const blk_q8_0 *)vy;
        for (int64_t ib = 0; ib < nb; ++ib) {
            int32_t sumi; uint16_t dx;
            if (xtype == OR_Q4_0) { const blk_q4_0 *x = (const blk_q4_0 *)vx + ib; oracle_blockdot_q4_0_q8_0(x, y + ib, 1, &sumi); dx = x->d; }
            else { const blk_q8_0 *x = (const blk_q8_0 *)vx + ib; oracle_blockdot_q8_0_q8_0(x, y + ib, 1, &sumi); dx = x->d; }
            sumf += sumi * h2f(dx) * h2f(y[ib].d);
        }
        return sumf;
    }
    if (xtype == OR_Q4_K) {
        const int64_t nb = n / QK_K; float sumf = 0, sumq = 0;
        const blk_q4_K *x = (const blk_q4_K *)vx; const blk_q8_K *y = (const blk_q8_K *)vy;
        for (int64_t i = 0; i < nb; ++i) {
            int32_t s[2]; oracle_blockdot_q4_K_q8_K(x + i, y + i, 1, s);
            const float d = h2f(x[i].d) * y[i].d, dmin = h2f(x[i].dmin) * y[i].d;
            sumq += d * s[0];
            sumf -= dmin * s[1];
        }
        return sumf + sumq;
    }
    if (xtype == OR_Q6_K) {   /* per super-block d_x * d_y * integer sum, fp32 in block order (ggml-quants.c:8677-8681) */
        const int64_t nb = n / QK_K; float sumf = 0;
        const blk_q6_K *x = (const blk_q6_K *)vx; const blk_q8_K *y = (const blk_q8_K *)vy;
        for (int64_t i = 0; i < nb; ++i) {
            int32_t si; oracle_blockdot_q6_K_q8_K(x + i, y + i, 1, &si);
            sumf += h2f(x[i].d) * y[i].d * si;
        }
        return sumf;
    }
    if (xtype == OR_Q5_0) {   /* ggml_vec_dot_q5_0_q8_0 generic branch, ggml-quants.c:4828-4848: sumf += (d_x*d_y) * sumi */
        const int64_t nb = n / QK; float sumf = 0;
        const blk_q5_0 *x = (const blk_q5_0 *)vx; const blk_q8_0 *y = (const blk_q8_0 *)vy;
        for (int64_t ib = 0; ib < nb; ++ib) {
            uint32_t qh; memcpy(&qh, x[ib].qh, 4);
            int sumi = 0;
            for (int j = 0; j < 16; ++j) {
                const uint8_t xh_0 = ((qh & (1u << (j + 0))) >> (j + 0)) << 4, xh_1 = ((qh & (1u << (j + 16))) >> (j + 12));
                const int x0 = ((x[ib].qs[j] & 0x0F) | xh_0) - 16, x1 = ((x[ib].qs[j] >> 4) | xh_1) - 16;
                sumi += x0 * y[ib].qs[j] + x1 * y[ib].qs[j + 16];
            }
            sumf += (h2f(x[ib].d) * h2f(y[ib].d)) * sumi;
        }
        return sumf;
    }
    if (xtype == OR_Q4_1) {   /* ggml_vec_dot_q4_1_q8_1 scalar, ggml-quants.c:4476-4492: sumf += (d_x*d_y)*sumi + m_x*s_y */
        const int64_t nb = n / QK; float sumf = 0;
        const blk_q4_1 *x = (const blk_q4_1 *)vx; const blk_q8_1 *y = (const blk_q8_1 *)vy;
        for (int64_t ib = 0; ib < nb; ++ib) {
            int sumi = 0;
            for (int j = 0; j < 16; ++j) sumi += (x[ib].qs[j] & 0x0F) * y[ib].qs[j] + (x[ib].qs[j] >> 4) * y[ib].qs[j + 16];
            sumf += (h2f(x[ib].d) * h2f(y[ib].d)) * sumi + h2f(x[ib].m) * h2f(y[ib].s);
        }
        return sumf;
    }
    if (xtype == OR_Q5_K) {   /* ggml_vec_dot_q5_K_q8_K generic branch, ggml-quants.c:7985-8032: per super-block d*sum_s sc_s*dot_s - dmin*sum_j m*bsums */
        const int64_t nb = n / QK_K; float sumq = 0, sumf = 0;
        const blk_q5_K *x = (const blk_q5_K *)vx; const blk_q8_K *y = (const blk_q8_K *)vy;
        for (int64_t i = 0; i < nb; ++i) {
            int32_t sq = 0, sm = 0;
            for (int s8 = 0; s8 < 8; ++s8) {
                uint8_t sc, m; get_scale_min_k4(s8, x[i].scales, &sc, &m);
                const uint8_t *ql = x[i].qs + 32 * (s8 / 2);
                int dot = 0;
                for (int l = 0; l < 32; ++l) {
                    const int q = ((s8 & 1) ? (ql[l] >> 4) : (ql[l] & 0xF)) + (((x[i].qh[l] >> s8) & 1) ? 16 : 0);
                    dot += q * y[i].qs[32 * s8 + l];
                }
                sq += sc * dot;
                sm += m * (y[i].bsums[2 * s8] + y[i].bsums[2 * s8 + 1]);
            }
            sumq += h2f(x[i].d) * y[i].d * sq;
            sumf -= h2f(x[i].dmin) * y[i].d * sm;
        }
        return sumf + sumq;
    }
    if (xtype == OR_IQ4_XS) {   /* ggml_vec_dot_iq4_xs_q8_K generic branch, ggml-quants.c:12100-12130: d4d8 * (ls - 32) * dot per 32-sub-block (integer sub-block sums here, one fp32 scale per super-block) */
        const int64_t nb = n / QK_K; float sumf = 0;
        const blk_iq4_xs *x = (const blk_iq4_xs *)vx; const blk_q8_K *y = (const blk_q8_K *)vy;
        for (int64_t i = 0; i < nb; ++i) {
            int32_t si = 0;
            for (int ib = 0; ib < 8; ++ib) {
                const int ls = ((x[i].scales_l[ib / 2] >> 4 * (ib % 2)) & 0xf) | (((x[i].scales_h >> 2 * ib) & 3) << 4);
                const uint8_t *qs = x[i].qs + 16 * ib;
                int dot = 0;
                for (int j = 0; j < 16; ++j) dot += kvalues_iq4nl[qs[j] & 0xf] * y[i].qs[32 * ib + j] + kvalues_iq4nl[qs[j] >> 4] * y[i].qs[32 * ib + 16 + j];
                si += (ls - 32) * dot;
            }
            sumf += h2f(x[i].d) * y[i].d * si;
        }
        return sumf;
    }
    return NAN;
}

/* ---------------------------------------------------------------- llamafile_sgemm */

static double dot_f16(const uint16_t *a, const void *b, int btype, long k) {
    double acc = 0;
    if (btype == OR_F16) { const uint16_t *bb = (const uint16_t *)b; for (long l = 0; l < k; ++l) acc += (double)(h2f(a[l]) * h2f(bb[l])); }
    else { const float *bb = (const float *)b; for (long l = 0; l < k; ++l) acc += (double)h2f(a[l]) * (double)bb[l]; }
    return acc;
}

/* Routing: llamafile/tinyblas_cpu_sgemm.inc:274-331 (entry) and :46-240 (impl), x86 AVX-512 build:
 *   Btype Q8_K  -> iqk (Q4_K here; k counts 256-blocks)              :288-292
 *   Btype Q8_0  -> iqk for Q4_0 (k counts 32-blocks)                 :293-303
 *   Q8_0 x Q8_0 -> tinyBLAS_Q0_AVX2                                  :183-192
 *   F16: B F32 only when n<=2, else B must be F16                    :121-136
 *   quantized A with F32 B -> false (WANT_QUANTIZATION)              :184-185, :204-205
 * C[ldc*j + i] = dot(A row i, B row j)                               tinyblas_cpu.h:600 */
int oracle_sgemm(long m, long n, long k, const void *A, long lda, const void *B, long ldb, void *C,
                 long ldc, int ith, int nth, int Atype, int Btype, int Ctype) {
    if (m < 0 || n < 0 || k < 0 || lda < k || ldb < k || ldc < m || nth <= 0 || ith >= nth) return 0;
    if (Ctype != OR_F32) return 0;
    float *c = (float *)C;
    if (Btype == OR_Q8_K) {
        if (Atype != OR_Q4_K && Atype != OR_Q6_K && Atype != OR_Q5_K && Atype != OR_IQ4_XS) return 0;   /* iqk_mul_mat.inc:1408-1463: the k-quants this build holds */
        const long rowa = (long)oracle_type_size(Atype);
        /* iqk ignores lda/ldb: rows are ggml_row_size apart (iqk_mul_mat.inc:191-199) */
        _Pragma("omp parallel for collapse(2)")
        for (long j = 0; j < n; ++j) for (long i = 0; i < m; ++i)
            c[ldc * j + i] = oracle_vec_dot(Atype, k * QK_K, (const char *)A + i * k * rowa, (const char *)B + j * k * 292);
        return 1;
    }
    if (Atype == OR_Q4_0 && Btype == OR_Q8_0) {
        _Pragma("omp parallel for collapse(2)")
        for (long j = 0; j < n; ++j) for (long i = 0; i < m; ++i)
            c[ldc * j + i] = oracle_vec_dot(OR_Q4_0, k * QK, (const char *)A + i * k * 18, (const char *)B + j * k * 34);
        return 1;
    }
    if ((Atype == OR_Q5_0 && Btype == OR_Q8_0) || (Atype == OR_Q4_1 && Btype == OR_Q8_1)) {   /* iqk_mul_mat.inc:1442-1451 */
        const long rowa = (long)oracle_type_size(Atype), rowb = (long)oracle_type_size(Btype);
        _Pragma("omp parallel for collapse(2)")
        for (long j = 0; j < n; ++j) for (long i = 0; i < m; ++i)
            c[ldc * j + i] = oracle_vec_dot(Atype, k * QK, (const char *)A + i * k * rowa, (const char *)B + j * k * rowb);
        return 1;
    }
    if (Atype == OR_Q8_0 && Btype == OR_Q8_0) {
        _Pragma("omp parallel for collapse(2)")
        for (long j = 0; j < n; ++j) for (long i = 0; i < m; ++i)
            c[ldc * j + i] = oracle_vec_dot(OR_Q8_0, k * QK, (const char *)A + i * lda * 34, (const char *)B + j * ldb * 34);
        return 1;
    }
    if (Atype == OR_F16) {
        if (Btype == OR_F32 && n > 2) return 0;
        if (Btype != OR_F32 && Btype != OR_F16) return 0;
        const size_t bsz = Btype == OR_F32 ? 4 : 2;
        _Pragma("omp parallel for collapse(2)")
        for (long j = 0; j < n; ++j) for (long i = 0; i < m; ++i)
            c[ldc * j + i] = (float)dot_f16((const uint16_t *)A + i * lda, (const char *)B + j * ldb * bsz, Btype, k);
        return 1;
    }
    if (Atype == OR_F32 && Btype == OR_F32) {
        const float *a = (const float *)A, *b = (const float *)B;
        _Pragma("omp parallel for collapse(2)")
        for (long j = 0; j < n; ++j) for (long i = 0; i < m; ++i) {
            double acc = 0; for (long l = 0; l < k; ++l) acc += (double)a[i * lda + l] * (double)b[j * ldb + l];
            c[ldc * j + i] = (float)acc;
        }
        return 1;
    }
    return 0;
}

/* ---------------------------------------------------------------- llamafile_mixmul */

#define ROW_ALIGN 64      /* llamafile/tinyblas_cpu.h:53-55 */
#define MATRIX_ALIGN 4096
#define MAX_ALIGN 4096

typedef struct {
    uintptr_t base;   /* params->wdata */
    char *wdata_;     /* aligned start */
    size_t allocated, wsize;
} mm_arena;

/* tinyblas_cpu_mixmul.inc:350-363 */
static char *mm_alloc(mm_arena *a, size_t align, size_t need) {
    size_t base = (a->allocated + align - 1) & -align;
    size_t toto = base + need;
    if (toto >= a->allocated && toto <= a->wsize) { a->allocated = toto; return a->wdata_ + base; }
    return NULL;
}

typedef struct { char *qthought; long ldq; int ok; mm_arena a; } mm_plan;

/* tinyblas_cpu_mixmul.inc:77-111 (constructor, allocate_shared_memory, get_allocated_bytes) */
static mm_plan mm_prepare(const oracle_tensor *w, const oracle_tensor *x, const oracle_tensor *plan, void *wdata, size_t wsize) {
    mm_plan p; memset(&p, 0, sizeof(p));
    const long cols = w->ne[0], experts = w->ne[2], thinkers = plan->ne[0], tasks = x->ne[1], tokens = x->ne[2];
    p.ldq = (cols * 2 + ROW_ALIGN - 1) & -ROW_ALIGN;
    p.a.base = (uintptr_t)wdata;
    p.a.wdata_ = (char *)(((uintptr_t)wdata + MAX_ALIGN - 1) & -(uintptr_t)MAX_ALIGN);
    p.a.wsize = wsize;
    p.qthought = mm_alloc(&p.a, MATRIX_ALIGN, (size_t)(tokens * tasks * p.ldq));
    p.ok = p.qthought != NULL
        && mm_alloc(&p.a, ROW_ALIGN, sizeof(uintptr_t) * experts * tokens * thinkers)
        && mm_alloc(&p.a, ROW_ALIGN, sizeof(uintptr_t) * experts * tokens * thinkers)
        && mm_alloc(&p.a, sizeof(long), sizeof(long) * experts);
    return p;
}

/* llamafile/tinyblas_cpu_mixmul_amd_avx.cpp:8-18 */
size_t oracle_mixmul_needs(const oracle_tensor *w, const oracle_tensor *x, const oracle_tensor *plan) {
    mm_plan p = mm_prepare(w, x, plan, (void *)0x1000, 0x7ffff000);
    return p.ok ? (size_t)((uintptr_t)p.a.wdata_ - p.a.base) + p.a.allocated : 0;
}

/* tinyblas_cpu_mixmul.inc:113-348. The per-expert pointer lists (build_row_pointers :297-320) are
 * an implementation device; semantically every (token, thinker) whose plan entry names a valid
 * expert gets result[token][thinker][:] = W[expert] . vec_dot_type(thought[token][thinker % tasks]).
 * Entries naming no expert in [0, experts) leave their result row untouched. */
int oracle_mixmul(const oracle_tensor *w, const oracle_tensor *x, const oracle_tensor *plan,
                  const oracle_tensor *out, void *wdata, size_t wsize) {
    mm_plan p = mm_prepare(w, x, plan, wdata, wsize);
    if (!p.ok) return 0;
    const long rows = w->ne[1], cols = w->ne[0], experts = w->ne[2], thinkers = plan->ne[0], tasks = x->ne[1], tokens = x->ne[2];
    if (!(tasks <= thinkers && thinkers <= experts && tokens == plan->ne[1] && rows == out->ne[0] && cols == x->ne[0] &&
          tokens == out->ne[2] && thinkers == out->ne[1] && plan->type == OR_I32)) return 0;   /* asserts :116-135 */
    if (w->nb[1] % oracle_type_size(w->type)) return 0;                 /* :138 */
    if (out->nb[0] != oracle_type_size(out->type) || x->nb[0] != oracle_type_size(x->type) ||
        w->nb[0] != oracle_type_size(w->type)) return 0;                /* :142-147 */
    if (out->type != OR_F32) return 0;
    int vdt, KN;
    switch (w->type) {                                                  /* mixmuler :162-269, AVX-512 build */
    case OR_F32: vdt = OR_F32; KN = 16; if (x->type != OR_F32) return 0; break;
    case OR_F16: vdt = OR_F16; KN = 16; if (x->type != OR_F32 && x->type != OR_F16) return 0; break;
    case OR_Q4_0: case OR_Q8_0: vdt = OR_Q8_0; KN = 32; if (x->type != OR_F32 && x->type != OR_Q8_0) return 0; break;
    default: return 0;
    }
    if (cols % KN) return 0;                                            /* :273 */
    const int need_q = x->type != vdt;
    if (need_q) {                                                       /* quantize_thought :322-348 */
        for (long t = 0; t < tokens; ++t) for (long task = 0; task < tasks; ++task) {
            const float *src = (const float *)((const char *)x->data + t * x->nb[2] + task * x->nb[1]);
            char *dst = p.qthought + t * tasks * p.ldq + task * p.ldq;
            if (vdt == OR_F16) oracle_fp32_to_fp16_row(src, (uint16_t *)dst, cols);
            else oracle_quantize_row_q8_0(src, dst, cols);
        }
    }
    for (long t = 0; t < tokens; ++t) for (long th = 0; th < thinkers; ++th) {
        const int32_t e = *(const int32_t *)((const char *)plan->data + t * plan->nb[1] + th * plan->nb[0]);
        if (e < 0 || e >= experts) continue;
        float *c = (float *)((char *)out->data + t * out->nb[2] + th * out->nb[1]);
        const char *b = need_q ? p.qthought + t * tasks * p.ldq + (th % tasks) * p.ldq
                               : (const char *)x->data + t * x->nb[2] + (th % tasks) * x->nb[1];
        const char *we = (const char *)w->data + e * w->nb[2];
        _Pragma("omp parallel for")
        for (long i = 0; i < rows; ++i) {
            const char *a = we + i * w->nb[1];
            if (w->type == OR_F16) c[i] = (float)dot_f16((const uint16_t *)a, b, OR_F16, cols);
            else if (w->type == OR_F32) { double acc = 0; for (long l = 0; l < cols; ++l) acc += (double)((const float *)a)[l] * ((const float *)b)[l]; c[i] = (float)acc; }
            else c[i] = oracle_vec_dot(w->type, cols, a, b);
        }
    }
    return 1;
}

/* llamafile/iqk_mul_mat.inc:204-221 with DataInfo::src1_row / dst_row (:84-101):
 * for y in [0,Ny): B row = (i1 % ne11 + i2*ne11), C row = C + i1*nb1/4 + i2*nb2/4 */
int oracle_mixmul_iqk(long Nx, long Ny, long ne00, int ne11, int typeA, const void *A, const void *B,
                      float *C, long nb1, long nb2, const void *row_mapping) {
    const int32_t *rm = (const int32_t *)row_mapping;
    int btype; size_t bs_a, bs_b; int blk;
    if (typeA == OR_Q4_K) { btype = OR_Q8_K; bs_a = 144; bs_b = 292; blk = 256; }
    else if (typeA == OR_Q6_K) { btype = OR_Q8_K; bs_a = 210; bs_b = 292; blk = 256; }
    else if (typeA == OR_Q4_0) { btype = OR_Q8_0; bs_a = 18; bs_b = 34; blk = 32; }
    else if (typeA == OR_Q5_K) { btype = OR_Q8_K; bs_a = 176; bs_b = 292; blk = 256; }
    else if (typeA == OR_IQ4_XS) { btype = OR_Q8_K; bs_a = 136; bs_b = 292; blk = 256; }
    else if (typeA == OR_Q5_0) { btype = OR_Q8_0; bs_a = 22; bs_b = 34; blk = 32; }
    else if (typeA == OR_Q4_1) { btype = OR_Q8_1; bs_a = 20; bs_b = 36; blk = 32; }
    else return 0;   /* other iqk types (:1408-1463) are outside the build's scope */
    (void)btype;
    if (ne00 % blk) return 0;
    const size_t row_a = (size_t)(ne00 / blk) * bs_a, row_b = (size_t)(ne00 / blk) * bs_b;
    for (long y = 0; y < Ny; ++y) {
        const int i1 = rm[2 * y], i2 = rm[2 * y + 1];
        const char *b = (const char *)B + (size_t)(i1 % ne11 + i2 * ne11) * row_b;
        float *c = C + i1 * (nb1 / 4) + i2 * (nb2 / 4);
        _Pragma("omp parallel for")
        for (long i = 0; i < Nx; ++i) c[i] = oracle_vec_dot(typeA, ne00, (const char *)A + i * row_a, b);
    }
    return 1;
}

/* ---------------------------------------------------------------------------------------------
 * The matmul's neighbours in the reference graph: llm_build_norm = ggml_rms_norm
 * then ggml_mul by the norm weight; llm_build_ffn (LLM_FFN_SILU, LLM_FFN_PAR) = ggml_silu(gate)
 * then ggml_mul by up (llama.cpp/src/llama.cpp:8099-8225).
 * --------------------------------------------------------------------------------------------- */

/* ggml_compute_forward_rms_norm_f32, llama.cpp/ggml/src/ggml.c:11850-11896: sum of x*x (float product) in
 * ggml_float = double, mean as float, scale = 1/sqrtf(mean + eps), y = x*scale (ggml_vec_scale_f32);
 * then ggml_mul: y[i] * w[i]. */
void oracle_rms_norm_mul(const float *x, const float *w, float eps, int64_t n, float *y) {
    double sum = 0.0;
    for (int64_t i = 0; i < n; ++i) sum += (double)(x[i] * x[i]);
    const float mean = (float)(sum / (double)n);
    const float scale = 1.0f / sqrtf(mean + eps);
    for (int64_t i = 0; i < n; ++i) {
        const float v = x[i] * scale;
        y[i] = v * w[i];
    }
}

/* ggml_silu_f32, ggml.c:2393-2395: x/(1.0f + expf(-x)) — the scalar definition; the reference's SIMD
 * ggml_v_silu (ggml.c:2432-2594) evaluates expf by a polynomial and differs from it by a few ulp, so this
 * function is pinned to the reference with a tolerance (tests/test_oracle_vs_reference.py), not bit for bit.
 * Then ggml_mul by up. */
void oracle_silu_mul(const float *gate, const float *up, int64_t n, float *y) {
    for (int64_t i = 0; i < n; ++i) {
        const float s = gate[i] / (1.0f + expf(-gate[i]));
        y[i] = s * up[i];
    }
}
