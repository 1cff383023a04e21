// Decode path (SURVEY.md §8 rows a9, a10, a11, a12 at n <= 8): skinny GEMV / GEMM that streams the
// resident quant planes from HBM with 128-bit coalesced loads and does int8 dp4a block dot products.
//
// Semantics follow the reference's block dot products:
//   Q4_0 x Q8_0 : sum_b d_x[b]*d_y[b] * sum_j (q4[j]-8)*q8[j]       ggml-quants.c:4192-4206 (ggml_vec_dot_q4_0_q8_0),
//                 served by iqk mul_mat_qX_0_q8_0_T<Q4_0_Unpacker>    llamafile/iqk_mul_mat.inc:1129-1201
//   Q8_0 x Q8_0 : sum_b d_x[b]*d_y[b] * sum_j q8x[j]*q8y[j]          tinyBLAS_Q0_AVX2, llamafile/tinyblas_cpu.h:934-993
//   Q4_K x Q8_K : sum_sb d*dy * sum_s sc_s*sum_j q4*q8 - dmin*dy * sum_s m_s*bsum_s   ggml-quants.c:7341-7397,
//                 served by iqk mul_mat_qX_K_q8_K_T<DequantizerQ4K>   llamafile/iqk_mul_mat.inc:601-643
//   F16         : fp32 FMA of widened fp16 (B fp32 when n<=2 else fp16)  llamafile/tinyblas_cpu_sgemm.inc:121-136
// The int32 block sums are exact; only the fp32 accumulation order differs from the AVX lanes.
//
// Mapping: one warp owns R weight rows and walks K; lane l takes 16-byte pieces l, l+32, ... of each
// row so every warp load is one contiguous 512-byte segment. Activations (quantized once per call by
// tb200_quantize.cu) are staged in shared memory, split so that lane-consecutive reads are
// bank-conflict free. Weights are read exactly once -> ld.global.nc.L1::no_allocate.
#include <stdlib.h>
#include <string.h>

#include "tb200_common.cuh"

namespace {

constexpr int kWarps = 8;   // warps per CTA

// ---- shared-memory activation staging ----------------------------------------------------------
// Q4_0: a 32-block's low-nibble elements are 0..15 and high-nibble elements 16..31, so the block's two
// 16-byte activation halves go to two planes: lo[t][b], hi[t][b]; lane b reads lo[b] and hi[b].
// Q8_0/F16: natural order. Q4_K: see gemv_q4_K.
// ---- Q4_0 ---------------------------------------------------------------------------------------
// U x R independent 16-byte weight loads are issued per lane before anything is consumed; the first batch is
// issued before the activations are even staged (weights are immutable, so it also precedes pdl_wait()), which
// folds the weight-fetch latency and the activation-fetch latency of these short kernels into one.
// One launch can serve up to kMaxMats weight matrices that share K and the activations (wq/wk/wv, gate/up): the
// grid is the concatenation of their row tiles. (Quantizing the fp32 activations inside every CTA was tried and
// measured slower at every shape — N=128256: 94.7 vs 48.1 us — so activations arrive pre-quantized.)
constexpr int kMaxMats = TB200_GEMV_MAX_MATS;
struct GemvQ4Args {
    int nmat;
    int world;                         // > 1: fused all-gather, lane p stores the row result into peer p's copy of the buffer
    const char *local_base;            // base of this rank's symmetric output buffer
    char *peer_base[TB200_MAX_PEERS];  // the same buffer on every rank (own entry included), peer-mapped
    int cta_begin[kMaxMats + 1];
    const void *qs[kMaxMats];
    const void *sc[kMaxMats];
    const void *d6[kMaxMats];           // Q6_K: fp16 super-block scale plane
    long row0[kMaxMats], nrows[kMaxMats];
    float *out[kMaxMats];
    long ldo[kMaxMats];
    long nb, nbc;
    const int8_t *aq; const int2 *ads; long a_stride;   // planar activations
    const tb200_gemv_item *items;
    const int *n_live;                 // grouped resident kernel: number of live items (they lead the table)
    int flags;                         // bit 0: the Q4_0 mma kernel pulls the step after next into L2 (TB200_MMA_PREFETCH=0 turns it off)
};

// Epilogue shared by the quantized decode kernels: warp-reduce each (row, token) and store. With world > 1 the
// store IS the all-gather: the CTA's kWarps*R row results are first collected in shared memory and then written as ONE
// contiguous segment (64 B at R = 2) into every rank's copy of the symmetric output buffer over NVLink (peer-mapped
// pointers), so no collective follows the kernel — only a flag barrier (tb200_peer.cu). (The first version let lane p
// of every warp store its 4-byte value to peer p: 16x more NVLink packets, and the barrier after the kernel waited
// for them.) Every warp of the CTA must call this (it contains a block barrier when world > 1).
template <int T, int R>
__device__ __forceinline__ void store_rows(const GemvQ4Args &A, int mat, float (&acc)[R][T], const int (&orow)[T], int ntok,
                                           long r_base, long nrows, int lane, bool active) {
    float *out = A.out[mat];
    const long ldo = A.ldo[mat];
    if (A.world <= 1) {
        if (!active) return;
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int t = 0; t < T; ++t) {
                const float v = warp_sum(acc[r][t]);
                if (lane == 0 && r_base + r < nrows && t < ntok) out[(long)orow[t] * ldo + r_base + r] = v;
            }
        return;
    }
    constexpr int kRowsPerCta = kWarps * R;
    __shared__ float s_out[T * kRowsPerCta];
    __shared__ int s_live[kWarps];                    // a warp without work this round (resident grids: the last round) must not store:
    const int warp = threadIdx.x >> 5;                // its r_base is the previous round's, and its sums are zero
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int t = 0; t < T; ++t) {
            const float v = warp_sum(acc[r][t]);
            if (lane == 0) s_out[t * kRowsPerCta + warp * R + r] = v;
        }
    if (lane == 0) s_live[warp] = active;
    __syncthreads();
    const long cta_base = r_base - (long)warp * R;   // first row of this CTA (relative to row0)
    const int j = threadIdx.x;                        // thread j < kRowsPerCta owns row j of the CTA: lanes write consecutive floats
    if (j < kRowsPerCta && s_live[j / R] && cta_base + j < nrows) {
#pragma unroll
        for (int t = 0; t < T; ++t)
            if (t < ntok) {
                const float v = s_out[t * kRowsPerCta + j];
                const long off = reinterpret_cast<const char *>(out + (long)orow[t] * ldo + cta_base + j) - A.local_base;
                for (int pr = 0; pr < A.world; ++pr) *reinterpret_cast<float *>(A.peer_base[pr] + off) = v;
            }
    }
    __syncthreads();                                  // resident grids call this once per round: s_out / s_live are rewritten next round
}

// (Unlike the K-quant kernels, this one loses with a shallower batch and four CTAs per SM: 3347 vs 3704 GB/s on the 8B step.)
template <int T, int R>
__global__ void __launch_bounds__(kWarps * 32) gemv_q4_0_kernel(const __grid_constant__ GemvQ4Args A) {
    constexpr int U = (T * R <= 4) ? 4 : 2;   // batch depth: blocks per lane per row in flight
    extern __shared__ __align__(16) uint8_t smem[];
    const long nb = A.nb, nbc = A.nbc;
    uint4 *s_lo = reinterpret_cast<uint4 *>(smem);
    uint4 *s_hi = s_lo + T * nbc;
    int2 *s_ds = reinterpret_cast<int2 *>(s_hi + T * nbc);
    __shared__ int s_arow[T];

    pdl_launch_dependents();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int mat = 0;
#pragma unroll
    for (int m = 1; m < kMaxMats; ++m)
        if (m < A.nmat && (int)blockIdx.x >= A.cta_begin[m]) mat = m;
    const long nrows = A.nrows[mat];
    const long r_base = ((long)((int)blockIdx.x - A.cta_begin[mat]) * kWarps + warp) * R;   // first row of this warp (relative to row0)
    const bool active = r_base < nrows;
    const uint4 *qs = (const uint4 *)A.qs[mat];
    const __half *sc = (const __half *)A.sc[mat];

    uint4 w[U][R];
    __half dh[U][R];
    const uint4 *wrow[R];
    const __half *srow[R];
    auto set_rows = [&](long my_row0) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const long row = my_row0 + min(r_base + r, nrows - 1);   // clamp: tail rows recompute the last row, never stored
            wrow[r] = qs + row * nb;
            srow[r] = sc + row * nb;
        }
    };
    auto load_batch = [&](long bbase, long cnt_end) {   // blocks bbase + lane + 32u, u < U, limited to < cnt_end (row-relative)
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long b = bbase + lane + 32 * u;
            if (b < cnt_end) {
#pragma unroll
                for (int r = 0; r < R; ++r) w[u][r] = ldg_stream(wrow[r] + b);
#pragma unroll
                for (int r = 0; r < R; ++r) dh[u][r] = srow[r][b];
            }
        }
    };

    // grouped mode (mixmul): this CTA column serves one work item = one expert's rows x up to T gathered tokens.
    // The item table is written by a predecessor kernel, so it is only read after pdl_wait().
    int ntok = T, orow[T];
#pragma unroll
    for (int t = 0; t < T; ++t) orow[t] = t;
    if (!A.items) {
        set_rows(A.row0[mat]);
        if (active) load_batch(0, min(nbc, nb));   // first batch in flight before the activations are touched
    }
    pdl_wait();
    if (A.items) {
        const tb200_gemv_item &it = A.items[blockIdx.y];
        ntok = it.ntok;
        if (ntok == 0) return;   // CTA-uniform
        set_rows(it.row0);
#pragma unroll
        for (int t = 0; t < T; ++t) orow[t] = it.out_row[min(t, ntok - 1)];
        if (threadIdx.x < T) s_arow[threadIdx.x] = it.a_row[min((int)threadIdx.x, ntok - 1)];
        if (active) load_batch(0, min(nbc, nb));
    } else if (threadIdx.x < T) s_arow[threadIdx.x] = threadIdx.x;
    __syncthreads();

    float acc[R][T];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int t = 0; t < T; ++t) acc[r][t] = 0.f;

    for (long b0 = 0; b0 < nb; b0 += nbc) {   // K chunks of nbc blocks (one chunk unless T*K is huge)
        const long cnt = min(nbc, nb - b0);
        if (b0) __syncthreads();
        const int nbc_i = (int)nbc;
        {
            // stage activations: 16-byte piece g of token t -> plane (g&1), index g>>1 (no div/mod: T is compile-time)
            const int cnt2 = (int)cnt * 2;
#pragma unroll
            for (int t = 0; t < T; ++t) {
                const uint4 *src = reinterpret_cast<const uint4 *>(A.aq + (long)s_arow[t] * A.a_stride + b0 * 32);
                const int2 *dsrc = A.ads + (long)s_arow[t] * (A.a_stride / QK) + b0;
                for (int g = threadIdx.x; g < cnt2; g += kWarps * 32) ((g & 1) ? s_hi : s_lo)[t * nbc_i + (g >> 1)] = src[g];
                for (int g = threadIdx.x; g < (int)cnt; g += kWarps * 32) s_ds[t * nbc_i + g] = dsrc[g];
            }
        }
        __syncthreads();
        if (!active) continue;
        for (long bb = 0; bb < cnt; bb += 32 * U) {
            if (b0 | bb) load_batch(b0 + bb, b0 + cnt);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const long b = bb + lane + 32 * u;   // chunk-relative
                if (b < cnt) {
                    // nibble planes: lo = elements 0..15, hi = elements 16..31 (values 0..15; the -8 is folded in below)
                    unsigned wl[R][4], wh[R][4];
                    float d[R];
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        d[r] = __half2float(dh[u][r]);
                        // high nibbles stay in place (= 16*q as unsigned bytes); their dp4a sum is shifted back once
                        wl[r][0] = w[u][r].x & 0x0f0f0f0fu, wh[r][0] = w[u][r].x & 0xf0f0f0f0u;
                        wl[r][1] = w[u][r].y & 0x0f0f0f0fu, wh[r][1] = w[u][r].y & 0xf0f0f0f0u;
                        wl[r][2] = w[u][r].z & 0x0f0f0f0fu, wh[r][2] = w[u][r].z & 0xf0f0f0f0u;
                        wl[r][3] = w[u][r].w & 0x0f0f0f0fu, wh[r][3] = w[u][r].w & 0xf0f0f0f0u;
                    }
#pragma unroll
                    for (int t = 0; t < T; ++t) {
                        const uint4 al = s_lo[t * nbc + b], ah = s_hi[t * nbc + b];
                        const int2 ds = s_ds[t * nbc + b];
                        const float dy = __int_as_float(ds.x);
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            int s = dp4a_us(wl[r][0], (int)al.x, -8 * ds.y);   // sum (q-8)*y = sum q*y - 8*sum y (exact in int32)
                            s = dp4a_us(wl[r][1], (int)al.y, s);
                            s = dp4a_us(wl[r][2], (int)al.z, s);
                            s = dp4a_us(wl[r][3], (int)al.w, s);
                            int sh = dp4a_us(wh[r][0], (int)ah.x, 0);
                            sh = dp4a_us(wh[r][1], (int)ah.y, sh);
                            sh = dp4a_us(wh[r][2], (int)ah.z, sh);
                            sh = dp4a_us(wh[r][3], (int)ah.w, sh);
                            acc[r][t] = fmaf(d[r] * dy, (float)(s + (sh >> 4)), acc[r][t]);   // sh is an exact multiple of 16
                        }
                    }
                }
            }
        }
    }
    store_rows<T, R>(A, mat, acc, orow, ntok, r_base, nrows, lane, active);
}

// ---- Q4_0, resident-grid form ----------------------------------------------------------------------
// Same arithmetic and per-warp work unit (R = 2 rows x one K batch) as gemv_q4_0_kernel, but the grid is exactly the
// number of CTAs the GPU can hold (3 per SM) and every warp walks the row pairs of the launch (all matrices
// concatenated) with a stride. The one-shot grid needs ceil(N/16) CTAs, i.e. 2.02 waves for N = 14336 and 18 for the
// vocabulary matrix: every wave boundary drains the memory pipe, the last wave is mostly empty, and each CTA stages
// the activations again (896 x 5 KB of L2 traffic for 33 MB of weights). Here they are staged once per CTA and all
// warps finish within one step of each other.
template <int T>
__global__ void __launch_bounds__(kWarps * 32, 3) gemv_q4_0_resident_kernel(const __grid_constant__ GemvQ4Args A) {
    constexpr int R = 2;
    constexpr int U = T <= 2 ? 4 : 2;
    extern __shared__ __align__(16) uint8_t smem[];
    const int nb = (int)A.nb;                      // whole K is staged (the launcher guarantees it fits)
    uint4 *s_lo = reinterpret_cast<uint4 *>(smem);
    uint4 *s_hi = s_lo + T * nb;
    int2 *s_ds = reinterpret_cast<int2 *>(s_hi + T * nb);

    pdl_launch_dependents();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nkb = (nb + 32 * U - 1) / (32 * U);  // K batches per row pair
    const long units = A.cta_begin[A.nmat];        // row pairs of all matrices (cta_begin[] holds unit offsets here)
    const long stride = (long)gridDim.x * kWarps;
    long unit = (long)blockIdx.x * kWarps + warp;

    int mat = 0;
    long r_base = 0, nrows = 0;
    const uint4 *wrow[R];
    const __half *srow[R];
    auto locate = [&](long u) {
        mat = 0;
#pragma unroll
        for (int m = 1; m < kMaxMats; ++m)
            if (m < A.nmat && u >= A.cta_begin[m]) mat = m;
        nrows = A.nrows[mat];
        r_base = (u - A.cta_begin[mat]) * R;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const long row = A.row0[mat] + min(r_base + r, nrows - 1);   // odd tail: the last row is recomputed, never stored
            wrow[r] = (const uint4 *)A.qs[mat] + row * nb;
            srow[r] = (const __half *)A.sc[mat] + row * nb;
        }
    };
    uint4 w[U][R];
    __half dh[U][R];
    auto load_batch = [&](int kbi) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int b = kbi * 32 * U + lane + 32 * u;
            if (b < nb) {
#pragma unroll
                for (int r = 0; r < R; ++r) w[u][r] = ldg_stream(wrow[r] + b);
#pragma unroll
                for (int r = 0; r < R; ++r) dh[u][r] = srow[r][b];
            }
        }
    };
    const bool any = unit < units;
    if (any) {
        locate(unit);
        load_batch(0);                              // immutable weights: in flight before the activations exist
    }
    pdl_wait();
#pragma unroll
    for (int t = 0; t < T; ++t) {
        const uint4 *src = reinterpret_cast<const uint4 *>(A.aq + (long)t * A.a_stride);
        const int2 *dsrc = A.ads + (long)t * (A.a_stride / QK);
        for (int g = threadIdx.x; g < nb * 2; g += kWarps * 32) ((g & 1) ? s_hi : s_lo)[t * nb + (g >> 1)] = src[g];
        for (int g = threadIdx.x; g < nb; g += kWarps * 32) s_ds[t * nb + g] = dsrc[g];
    }
    __syncthreads();

    int orow[T];
#pragma unroll
    for (int t = 0; t < T; ++t) orow[t] = t;
    const bool peers = A.world > 1;                 // the peer-store epilogue contains a block barrier: every warp must reach it
    const long rounds = (units + stride - 1) / stride;
    for (long it = 0; it < rounds; ++it, unit += stride) {
        const bool active = unit < units;
        float acc[R][T];
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int t = 0; t < T; ++t) acc[r][t] = 0.f;
        if (active) {
            if (it) locate(unit);
            for (int kb = 0; kb < nkb; ++kb) {
                if (it | kb) load_batch(kb);
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int b = kb * 32 * U + lane + 32 * u;
                    if (b < nb) {
                        unsigned wl[R][4], wh[R][4];
                        float d[R];
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            d[r] = __half2float(dh[u][r]);
                            wl[r][0] = w[u][r].x & 0x0f0f0f0fu, wh[r][0] = w[u][r].x & 0xf0f0f0f0u;
                            wl[r][1] = w[u][r].y & 0x0f0f0f0fu, wh[r][1] = w[u][r].y & 0xf0f0f0f0u;
                            wl[r][2] = w[u][r].z & 0x0f0f0f0fu, wh[r][2] = w[u][r].z & 0xf0f0f0f0u;
                            wl[r][3] = w[u][r].w & 0x0f0f0f0fu, wh[r][3] = w[u][r].w & 0xf0f0f0f0u;
                        }
#pragma unroll
                        for (int t = 0; t < T; ++t) {
                            const uint4 al = s_lo[t * nb + b], ah = s_hi[t * nb + b];
                            const int2 ds = s_ds[t * nb + b];
                            const float dy = __int_as_float(ds.x);
#pragma unroll
                            for (int r = 0; r < R; ++r) {
                                int sl = dp4a_us(wl[r][0], (int)al.x, -8 * ds.y);   // sum (q-8)*y = sum q*y - 8*sum y (exact in int32)
                                sl = dp4a_us(wl[r][1], (int)al.y, sl);
                                sl = dp4a_us(wl[r][2], (int)al.z, sl);
                                sl = dp4a_us(wl[r][3], (int)al.w, sl);
                                int sh = dp4a_us(wh[r][0], (int)ah.x, 0);
                                sh = dp4a_us(wh[r][1], (int)ah.y, sh);
                                sh = dp4a_us(wh[r][2], (int)ah.z, sh);
                                sh = dp4a_us(wh[r][3], (int)ah.w, sh);
                                acc[r][t] = fmaf(d[r] * dy, (float)(sl + (sh >> 4)), acc[r][t]);   // sh is an exact multiple of 16
                            }
                        }
                    }
                }
            }
        }
        if (active || peers) store_rows<T, R>(A, mat, acc, orow, T, r_base, nrows, lane, active);
    }
}

// ---- Q4_0, grouped resident form (mixmul decode, more routed pairs than experts) ---------------------------------
// One item = all rows of one expert against up to T gathered tokens. The one-shot grouped launch gives every CTA 16
// weight rows (36 KB at K = 4096) and makes it stage the item's T activation rows first (20 KB at T = 4), and its token
// loop always runs T wide. Here the grid is the CTAs the GPU holds; each CTA serves one of the L live items (L written by
// the item builder, read after pdl_wait), dealt out by cost, and walks that expert's row pairs with the stride of the CTAs
// sharing the item, so the activations are staged once per CTA and amortised over ~25 row pairs.
//
// Token loops are bounded by the item's real token count (CTA-uniform), so a half-empty group costs half the dp4a work.
// Activations are staged per K block as one record {T x low halves | T x high halves | T x {d, sum} | 16 B pad}: every
// token's operands sit at a compile-time offset from ONE address per block, and the pad makes the record stride 4 (mod 8)
// words, which keeps the 128-bit lane accesses conflict-free. The kernel is compiled for two CTAs per SM (128 registers)
// so that the nibble masks of a block stay in registers across its tokens: with the 80 registers of three CTAs per SM and
// planar staging, ptxas re-derived the masks and three staging addresses inside every guarded token block (ncu, batch 32:
// 44 instructions per 32-element block and token - 19 % LOP3, 18 % LEA/IMAD - against 8 dp4a; issue-bound at 290 us).
// Compiling the body once per token count inside a switch was tried: each body alone is spill-free, eight together spill 2 KB.
template <int NT>
__device__ __forceinline__ void grouped_item_q4_0(const GemvQ4Args &A, const tb200_gemv_item &it, int slot, int sharers, uint8_t *smem) {
    constexpr int R = 2;
    constexpr int U = 4;
    constexpr int kRec = NT * 40 + 16;
    const int ntok = it.ntok;                      // 1..NT, CTA-uniform
    const int nb = (int)A.nb;                      // whole K is staged (the launcher guarantees it fits)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long row0 = it.row0, nrows = A.nrows[0];
    const int nkb = (nb + 32 * U - 1) / (32 * U);  // K batches per row pair
    const long units = (nrows + 1) / 2;
    const long stride = (long)sharers * kWarps;
    long unit = (long)slot * kWarps + warp;
    const uint4 *qs = (const uint4 *)A.qs[0];
    const __half *sc = (const __half *)A.sc[0];
    const uint4 *wrow[R];
    const __half *srow[R];
    long r_base = 0;
    auto locate = [&](long u) {
        r_base = u * R;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const long row = row0 + min(r_base + r, nrows - 1);   // odd tail: the last row is recomputed, never stored
            wrow[r] = qs + row * nb;
            srow[r] = sc + row * nb;
        }
    };
    uint4 w[U][R];
    __half dh[U][R];
    auto load_batch = [&](int kbi) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int b = kbi * 32 * U + lane + 32 * u;
            if (b < nb) {
#pragma unroll
                for (int r = 0; r < R; ++r) w[u][r] = ldg_stream(wrow[r] + b);
#pragma unroll
                for (int r = 0; r < R; ++r) dh[u][r] = srow[r][b];
            }
        }
    };
    if (unit < units) {
        locate(unit);
        load_batch(0);
    }
    for (int t = 0; t < ntok; ++t) {
        const uint4 *src = reinterpret_cast<const uint4 *>(A.aq + (long)it.a_row[t] * A.a_stride);
        const int2 *dsrc = A.ads + (long)it.a_row[t] * (A.a_stride / QK);
        for (int g = threadIdx.x; g < nb * 2; g += kWarps * 32)
            *reinterpret_cast<uint4 *>(smem + (g >> 1) * kRec + (g & 1) * (NT * 16) + t * 16) = src[g];
        for (int g = threadIdx.x; g < nb; g += kWarps * 32) *reinterpret_cast<int2 *>(smem + g * kRec + NT * 32 + t * 8) = dsrc[g];
    }
    __syncthreads();

    float *out = A.out[0];
    const long ldo = A.ldo[0];
    for (bool first = true; unit < units; unit += stride, first = false) {
        float acc[R][NT];
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int t = 0; t < NT; ++t) acc[r][t] = 0.f;
        if (!first) locate(unit);
        for (int kb = 0; kb < nkb; ++kb) {
            if (!first || kb) load_batch(kb);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int b = kb * 32 * U + lane + 32 * u;
                if (b < nb) {
                    const uint8_t *rec = smem + b * kRec;
                    unsigned wl[R][4], wh[R][4];
                    float d[R];
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        d[r] = __half2float(dh[u][r]);
                        wl[r][0] = w[u][r].x & 0x0f0f0f0fu, wh[r][0] = w[u][r].x & 0xf0f0f0f0u;
                        wl[r][1] = w[u][r].y & 0x0f0f0f0fu, wh[r][1] = w[u][r].y & 0xf0f0f0f0u;
                        wl[r][2] = w[u][r].z & 0x0f0f0f0fu, wh[r][2] = w[u][r].z & 0xf0f0f0f0u;
                        wl[r][3] = w[u][r].w & 0x0f0f0f0fu, wh[r][3] = w[u][r].w & 0xf0f0f0f0u;
                    }
#pragma unroll
                    for (int t = 0; t < NT; ++t) {
                        if (t >= ntok) break;
                        const uint4 al = *reinterpret_cast<const uint4 *>(rec + t * 16);
                        const uint4 ah = *reinterpret_cast<const uint4 *>(rec + NT * 16 + t * 16);
                        const int2 ds = *reinterpret_cast<const int2 *>(rec + NT * 32 + t * 8);
                        const float dy = __int_as_float(ds.x);
                        const int bias = -8 * ds.y;
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            int sl = dp4a_us(wl[r][0], (int)al.x, bias);   // sum (q-8)*y = sum q*y - 8*sum y (exact in int32)
                            sl = dp4a_us(wl[r][1], (int)al.y, sl);
                            sl = dp4a_us(wl[r][2], (int)al.z, sl);
                            sl = dp4a_us(wl[r][3], (int)al.w, sl);
                            int sh = dp4a_us(wh[r][0], (int)ah.x, 0);
                            sh = dp4a_us(wh[r][1], (int)ah.y, sh);
                            sh = dp4a_us(wh[r][2], (int)ah.z, sh);
                            sh = dp4a_us(wh[r][3], (int)ah.w, sh);
                            acc[r][t] = fmaf(d[r] * dy, (float)(sl + (sh >> 4)), acc[r][t]);   // sh is an exact multiple of 16
                        }
                    }
                }
            }
        }
#pragma unroll
        for (int t = 0; t < NT; ++t) {
            if (t >= ntok) break;
            const long o = (long)it.out_row[t] * ldo + r_base;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float v = warp_sum(acc[r][t]);
                if (lane == 0 && r_base + r < nrows) out[o + r] = v;
            }
        }
    }
}

template <int T>
__global__ void __launch_bounds__(kWarps * 32, 2) gemv_q4_0_grouped_kernel(const __grid_constant__ GemvQ4Args A) {
    extern __shared__ __align__(16) uint8_t smem[];
    pdl_launch_dependents();
    pdl_wait();                                    // item table and live count come from the predecessor
    const int live = *A.n_live;
    if (live <= 0) return;
    // Which item this CTA serves. An item costs about (ntok + 1) units (one pass over the expert's weights + ntok dot products
    // per block), and the kernel ends with its slowest item, so CTAs are dealt out in proportion to that cost: one per item,
    // the rest by cumulative cost. With more items than half the grid every CTA walks the items c, c + grid, ... alone.
    constexpr int kMaxLive = kWarps * 32;
    __shared__ int s_cum[kMaxLive + 1], s_pick[3];
    const bool dealt = 2 * live <= (int)gridDim.x && live <= kMaxLive;
    if (dealt) {
        if ((int)threadIdx.x < live) s_cum[threadIdx.x + 1] = A.items[threadIdx.x].ntok + 1;
        __syncthreads();
        if (threadIdx.x == 0) {
            s_cum[0] = 0;
            for (int i = 0; i < live; ++i) s_cum[i + 1] += s_cum[i];
        }
        __syncthreads();
        const long spare = (long)gridDim.x - live, total = s_cum[live];
        if ((int)threadIdx.x < live) {
            const int lo = (int)threadIdx.x + (int)(s_cum[threadIdx.x] * spare / total), hi = (int)threadIdx.x + 1 + (int)(s_cum[threadIdx.x + 1] * spare / total);
            if ((int)blockIdx.x >= lo && (int)blockIdx.x < hi) s_pick[0] = threadIdx.x, s_pick[1] = (int)blockIdx.x - lo, s_pick[2] = hi - lo;
        }
        __syncthreads();
    }
    for (int item = dealt ? s_pick[0] : (int)blockIdx.x; item < live; item += dealt ? live : (int)gridDim.x) {
        const int slot = dealt ? s_pick[1] : 0, sharers = dealt ? s_pick[2] : 1;
        const tb200_gemv_item &it = A.items[item];
        grouped_item_q4_0<T>(A, it, slot, sharers, smem);
        __syncthreads();   // the staged activations are replaced by the next item's
    }
}

// ---- Q4_0, skinny GEMM on the integer tensor-core path (2..8 tokens) ---------------------------------------------
// With T tokens the dp4a kernels spend 8 dp4a + ~12 other instructions per 32-element block, row and token: at T = 8
// the vocabulary matrix took 209 us (1.4 TB/s) against 46 us at T = 1 - the kernel is issue-bound, not HBM-bound. One
// warp-level mma.sync.m16n8k32 (u8 x s8 -> s32) does a whole Q4_0 block of 16 rows against 8 tokens exactly as the block dot
// product defines it: K = 32 is ONE block, so the int32 result is the block's sum_j q[j]*y[j] (ggml-quants.c:4192-4206) and
// is scaled per block in fp32 like the reference. The -8 of (q - 8) enters as the accumulator's initial value -8*sum(y),
// so the int32 the scale multiplies is exact.
//   A fragment: lane (g = lane/4, c = lane%4) needs word c of the packed block of rows g and g+8: its low nibbles are
//   elements 4c..4c+3 (a0/a1), its high nibbles elements 16+4c..16+4c+3 (a2/a3) - exactly the K slices the fragment wants,
//   so unpacking is two ANDs and a shift. The weights go from HBM to registers with no staging: lane (g, c) loads the whole
//   16-byte block b0+c of its rows (a warp load touches 8 rows x 64 contiguous bytes) and the four lanes of a group exchange
//   words with three shuffles (4x4 transpose). (First version: 4-byte loads of word c straight from HBM - every warp load
//   touched 8 cache lines for 128 bytes and the L1 tag stage capped the kernel at 2.1 TB/s.)
//   B fragment: the activations of token g, same K slices, staged once per CTA as [block][token][c]{b0, b1} (one
//   conflict-free 8-byte read per mma); {d_y, -8*sum y} per block and token sit beside them.
// A CTA owns one 16-row tile at a time and its 8 warps split K, so even a 4096-row matrix fills the GPU (256 tiles); the
// eight partial 16x8 tiles meet in shared memory. The grid is resident (CTAs walk the tiles with a stride) and every lane
// keeps the next batch of blocks in flight while it consumes one.
__device__ __forceinline__ void mma_u8s8_16832(int (&d)[4], unsigned a0, unsigned a1, unsigned a2, unsigned a3, unsigned b0, unsigned b1, int c0, int c1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%11,%10,%11};"
                 : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(c0), "r"(c1));
}
// lane c of a 4-lane group holds x = the four words of block c; on return x[j] = word c of block j
__device__ __forceinline__ void quad_transpose(uint4 &x, int c, int lane) {
    unsigned z0 = x.x, z1 = x.y, z2 = x.z, z3 = x.w, t;
    if (c & 1) t = z0, z0 = z1, z1 = z2, z2 = z3, z3 = t;             // z[i] = x[(i + c) & 3]
    if (c & 2) t = z0, z0 = z2, z2 = t, t = z1, z1 = z3, z3 = t;
    const int base = lane & ~3;
    const unsigned w1 = __shfl_sync(0xffffffffu, z1, base | ((c - 1) & 3));   // round r: word c of block (c - r) & 3
    const unsigned w2 = __shfl_sync(0xffffffffu, z2, base | ((c - 2) & 3));
    const unsigned w3 = __shfl_sync(0xffffffffu, z3, base | ((c - 3) & 3));
    unsigned v0 = z0, v1 = w3, v2 = w2, v3 = w1;                        // v[i] = word c of block (c + i) & 3
    if (c & 1) t = v3, v3 = v2, v2 = v1, v1 = v0, v0 = t;             // y[j] = v[(j - c) & 3]
    if (c & 2) t = v0, v0 = v2, v2 = t, t = v1, v1 = v3, v3 = t;
    x = make_uint4(v0, v1, v2, v3);
}

// PAIR (TB200_MMA_PAIR=1, experimental): 8-byte loads instead - lanes c < 2 hold the halves of one block, lanes c >= 2 those of the
// next, and one shuffle with lane c^2 + three selects per two mma replace the 4x4 transpose (lane c then carries word {0,2,1,3}[c]).
__device__ __forceinline__ uint2 ldg_stream64(const uint2 *p) {
    uint2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
    return r;
}
template <int T, int KS, bool PAIR = false>   // KS = warps sharing a tile's K (8: small matrices, every tile's partials meet in shared memory; 1: a warp owns its tile)
__global__ void __launch_bounds__(kWarps * 32, 3) gemm_q4_0_mma_kernel(const __grid_constant__ GemvQ4Args A) {
    constexpr int Q = 2;                           // quads (4 blocks) per lane and row in flight per batch
    extern __shared__ __align__(16) uint8_t smem[];
    const int nb = (int)A.nb;                      // nb % 32 == 0 (launcher): every warp's K slice is whole quads
    uint2 *s_b = reinterpret_cast<uint2 *>(smem);                        // [nb][T][4] {b0, b1}
    int4 *s_ds = reinterpret_cast<int4 *>(smem + (size_t)nb * T * 32);   // [nb][4] {d_y(2c), bias(2c), d_y(2c+1), bias(2c+1)}
    __shared__ float4 s_red[KS > 1 ? 2 : 1][KS > 1 ? kWarps : 1][32];

    pdl_launch_dependents();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, c = lane & 3;
    const long nrows = A.nrows[0];
    const long tiles = (nrows + 15) / 16;
    const int slice = nb / KS;                     // blocks of K per warp
    const int b_lo = KS > 1 ? warp * slice : 0;
    const int nbat = (slice + 4 * Q - 1) / (4 * Q);
    const long tile0 = KS > 1 ? (long)blockIdx.x : (long)blockIdx.x * kWarps + warp, tstride = KS > 1 ? (long)gridDim.x : (long)gridDim.x * kWarps;
    const long my_tiles = tile0 < tiles ? (tiles - tile0 + tstride - 1) / tstride : 0;
    const long steps = my_tiles * nbat;
    const uint4 *qs = (const uint4 *)A.qs[0];
    const __half *sc = (const __half *)A.sc[0];

    // prefetch side: step -> (tile, K batch); lane (g, c) loads block (quad base + c) of rows g and g+8. The four row
    // pointers are rebuilt once per tile and advanced by a constant per batch.
    long pf_tile = tile0;
    int pf_bat = 0;
    const uint4 *pq0 = nullptr, *pq1 = nullptr;
    const uint2 *ps0 = nullptr, *ps1 = nullptr;
    auto load_step = [&](uint4 (&w)[Q][2], uint2 (&dq)[Q][2]) {
        if (pf_bat == 0) {
            const long r0 = A.row0[0] + min(pf_tile * 16 + g, nrows - 1), r1 = A.row0[0] + min(pf_tile * 16 + g + 8, nrows - 1);
            pq0 = qs + r0 * nb + b_lo + (PAIR ? 0 : c), pq1 = qs + r1 * nb + b_lo + (PAIR ? 0 : c);
            ps0 = reinterpret_cast<const uint2 *>(sc + r0 * nb + b_lo), ps1 = reinterpret_cast<const uint2 *>(sc + r1 * nb + b_lo);
        }
#pragma unroll
        for (int q = 0; q < Q; ++q)
            if (pf_bat * 4 * Q + q * 4 < slice) {   // slice-relative, warp-uniform
                if constexpr (PAIR) {
                    const uint2 *h0p = reinterpret_cast<const uint2 *>(pq0 + q * 4) + c, *h1p = reinterpret_cast<const uint2 *>(pq1 + q * 4) + c;
                    const uint2 a0 = ldg_stream64(h0p), b0 = ldg_stream64(h0p + 4), a1 = ldg_stream64(h1p), b1 = ldg_stream64(h1p + 4);
                    w[q][0] = make_uint4(a0.x, a0.y, b0.x, b0.y), w[q][1] = make_uint4(a1.x, a1.y, b1.x, b1.y);
                } else {
                    w[q][0] = ldg_stream(pq0 + q * 4), w[q][1] = ldg_stream(pq1 + q * 4);
                }
                dq[q][0] = __ldg(ps0 + q), dq[q][1] = __ldg(ps1 + q);
            }
        pq0 += 4 * Q, pq1 += 4 * Q, ps0 += Q, ps1 += Q;
        if (++pf_bat == nbat) pf_bat = 0, pf_tile += tstride;
        if ((A.flags & 1) && lane < 16 && pf_tile < tiles) {   // the step after this one: one 128-byte line per row into L2
            const long r = A.row0[0] + min(pf_tile * 16 + lane, nrows - 1);
            asm volatile("prefetch.global.L2 [%0];" ::"l"(qs + r * nb + b_lo + pf_bat * 4 * Q));
        }
    };
    uint4 wA[Q][2], wB[Q][2];
    uint2 dA[Q][2], dB[Q][2];
    if (steps > 0) load_step(wA, dA);              // immutable weights: in flight before the activations exist

    pdl_wait();
    // stage the activations in fragment order: source word w of (token, block) -> [block][token][w & 3] slot (w >> 2)
    {
        unsigned *sb = reinterpret_cast<unsigned *>(s_b);
        const int words = nb * 8;
#pragma unroll
        for (int t = 0; t < T; ++t) {
            const unsigned *src = reinterpret_cast<const unsigned *>(A.aq + (long)t * A.a_stride);
            for (int i = threadIdx.x; i < words; i += kWarps * 32) {
                const int b = i >> 3, w = i & 7;
                const int cc = PAIR ? (((w & 1) << 1) | ((w >> 1) & 1)) : (w & 3);   // PAIR: lane c carries word {0,2,1,3}[c]
                sb[((b * T + t) * 4 + cc) * 2 + (w >> 2)] = src[i];
            }
        }
        int *sd = reinterpret_cast<int *>(s_ds);
        for (int i = threadIdx.x; i < nb * 8; i += kWarps * 32) {
            const int b = i >> 3, t = i & 7;
            int2 v = make_int2(0, 0);
            if (t < T) v = A.ads[(long)t * (A.a_stride / QK) + b];
            sd[(b * 8 + t) * 2] = v.x, sd[(b * 8 + t) * 2 + 1] = -8 * v.y;
        }
    }
    __syncthreads();

    const int tb = g < T ? g : T - 1;              // B columns beyond T repeat the last token; their results are never stored
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    long cu_tile = tile0;
    int cu_bat = 0, parity = 0;
    float *out = A.out[0];
    const long ldo = A.ldo[0];
    auto compute_step = [&](uint4 (&w)[Q][2], const uint2 (&dq)[Q][2]) {
#pragma unroll
        for (int q = 0; q < Q; ++q) {
            const int bq = cu_bat * 4 * Q + q * 4;
            if (bq < slice) {                      // warp-uniform
                if constexpr (PAIR) {
#pragma unroll
                    for (int r = 0; r < 2; ++r) {
                        const uint4 x = w[q][r];
                        const unsigned ra = __shfl_xor_sync(0xffffffffu, c < 2 ? x.y : x.x, 2), rb = __shfl_xor_sync(0xffffffffu, c < 2 ? x.w : x.z, 2);
                        w[q][r] = make_uint4(c < 2 ? x.x : ra, c < 2 ? ra : x.y, c < 2 ? x.z : rb, c < 2 ? rb : x.w);
                    }
                } else {
                    quad_transpose(w[q][0], c, lane);
                    quad_transpose(w[q][1], c, lane);
                }
                const unsigned w0[4] = {w[q][0].x, w[q][0].y, w[q][0].z, w[q][0].w}, w1[4] = {w[q][1].x, w[q][1].y, w[q][1].z, w[q][1].w};
                const unsigned h0[2] = {dq[q][0].x, dq[q][0].y}, h1[2] = {dq[q][1].x, dq[q][1].y};
                const uint2 *pb = s_b + ((b_lo + bq) * T + tb) * 4 + c;   // one address per quad, compile-time offsets per block
                const int4 *pd = s_ds + (b_lo + bq) * 4 + c;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint2 bf = pb[j * T * 4];
                    const int4 ds = pd[j * 4];
                    int d[4];
                    mma_u8s8_16832(d, w0[j] & 0x0f0f0f0fu, w1[j] & 0x0f0f0f0fu, (w0[j] >> 4) & 0x0f0f0f0fu, (w1[j] >> 4) & 0x0f0f0f0fu, bf.x, bf.y, ds.y, ds.w);
                    const float dw0 = __half2float(__ushort_as_half((unsigned short)((j & 1) ? (h0[j >> 1] >> 16) : (h0[j >> 1] & 0xffffu))));
                    const float dw1 = __half2float(__ushort_as_half((unsigned short)((j & 1) ? (h1[j >> 1] >> 16) : (h1[j >> 1] & 0xffffu))));
                    const float y0 = __int_as_float(ds.x), y1 = __int_as_float(ds.z);
                    acc[0] = fmaf(dw0 * y0, (float)d[0], acc[0]);
                    acc[1] = fmaf(dw0 * y1, (float)d[1], acc[1]);
                    acc[2] = fmaf(dw1 * y0, (float)d[2], acc[2]);
                    acc[3] = fmaf(dw1 * y1, (float)d[3], acc[3]);
                }
            }
        }
        if (++cu_bat == nbat) {
            if constexpr (KS > 1) {   // this warp's K slice of the tile is done: the eight partial tiles meet in shared memory
                s_red[parity][warp][lane] = make_float4(acc[0], acc[1], acc[2], acc[3]);
                __syncthreads();
                if (threadIdx.x < 128) {
                    const float *p = reinterpret_cast<const float *>(&s_red[parity][0][0]) + threadIdx.x;
                    float v = 0.f;
#pragma unroll
                    for (int w8 = 0; w8 < kWarps; ++w8) v += p[w8 * 128];
                    const int ln = threadIdx.x >> 2, i = threadIdx.x & 3;   // lane (g, c) element i: row g + 8*(i>>1), token 2c + (i&1)
                    const long r = cu_tile * 16 + (ln >> 2) + 8 * (i >> 1);
                    const int tok = 2 * (ln & 3) + (i & 1);
                    if (tok < T && r < nrows) out[(long)tok * ldo + r] = v;
                }
                parity ^= 1;   // the next tile's partials go to the other buffer, so one barrier per tile is enough
            } else {                  // lane (g, c) holds rows g, g+8 x tokens 2c, 2c+1, complete
                const long r = cu_tile * 16 + g;
                if (2 * c < T) {
                    if (r < nrows) out[(long)(2 * c) * ldo + r] = acc[0];
                    if (r + 8 < nrows) out[(long)(2 * c) * ldo + r + 8] = acc[2];
                }
                if (2 * c + 1 < T) {
                    if (r < nrows) out[(long)(2 * c + 1) * ldo + r] = acc[1];
                    if (r + 8 < nrows) out[(long)(2 * c + 1) * ldo + r + 8] = acc[3];
                }
            }
            acc[0] = acc[1] = acc[2] = acc[3] = 0.f;
            cu_bat = 0, cu_tile += tstride;
        }
    };
    for (long s = 0; s < steps; s += 2) {
        if (s + 1 < steps) load_step(wB, dB);
        compute_step(wA, dA);
        if (s + 1 >= steps) break;
        if (s + 2 < steps) load_step(wA, dA);
        compute_step(wB, dB);
    }
}

// ---- Q8_0, skinny GEMM on the integer tensor-core path (4..8 tokens) --------------------------------------------
// Same plan as gemm_q4_0_mma_kernel for tinyBLAS_Q0_AVX2's Q8_0 x Q8_0 block product (llamafile/tinyblas_cpu.h:934-993):
// one mma.sync.m16n8k32 (s8 x s8 -> s32) per 32-byte block of 16 rows x 8 tokens, fp32 scaling per block. A block is two
// 16-byte units; lane (g, c) loads unit c of a PAIR of blocks of its rows (a warp load = 8 rows x 64 contiguous bytes), lanes
// c and c^2 swap eight bytes, and every lane then holds eight bytes of the first and eight of the second block: lane c
// carries bytes {0,16,8,24}[c]..+7 of either block. The dot product is order-free inside a block, so the B fragment (the
// token's activations) is simply staged with the same byte order.
__device__ __forceinline__ void mma_s8s8_16832(int (&d)[4], unsigned a0, unsigned a1, unsigned a2, unsigned a3, unsigned b0, unsigned b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
                 : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "r"(0));
}

template <int T>
__global__ void __launch_bounds__(kWarps * 32, 3) gemm_q8_0_mma_kernel(const uint4 *__restrict__ qs, const __half *__restrict__ sc, long nb, long row0, long nrows,
                                                                       const int8_t *__restrict__ aq, const int2 *__restrict__ ads, long a_stride,
                                                                       float *__restrict__ out, long ldo) {
    constexpr int P = 2;                           // block pairs per lane and row in flight per batch (4 blocks)
    extern __shared__ __align__(16) uint8_t smem[];
    uint2 *s_b = reinterpret_cast<uint2 *>(smem);                                  // [nb][T][4] {b0, b1}
    float2 *s_dy = reinterpret_cast<float2 *>(smem + (size_t)nb * T * 32);         // [nb][4] {d_y(2c), d_y(2c+1)}
    __shared__ float4 s_red[2][kWarps][32];

    pdl_launch_dependents();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, c = lane & 3;
    const long tiles = (nrows + 15) / 16;
    const int slice = (int)nb / kWarps;            // blocks of K per warp; nb % 32 == 0 (launcher)
    const int b_lo = warp * slice;
    const int nbat = slice / (2 * P);
    const long my_tiles = (long)blockIdx.x < tiles ? (tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const long steps = my_tiles * nbat;

    long pf_tile = blockIdx.x;
    int pf_bat = 0;
    const uint4 *pq0 = nullptr, *pq1 = nullptr;
    const uint2 *ps0 = nullptr, *ps1 = nullptr;
    auto load_step = [&](uint4 (&w)[P][2], uint2 (&dq)[2]) {
        if (pf_bat == 0) {
            const long r0 = row0 + min(pf_tile * 16 + g, nrows - 1), r1 = row0 + min(pf_tile * 16 + g + 8, nrows - 1);
            pq0 = qs + (r0 * nb + b_lo) * 2 + c, pq1 = qs + (r1 * nb + b_lo) * 2 + c;
            ps0 = reinterpret_cast<const uint2 *>(sc + r0 * nb + b_lo), ps1 = reinterpret_cast<const uint2 *>(sc + r1 * nb + b_lo);
        }
#pragma unroll
        for (int p = 0; p < P; ++p) w[p][0] = ldg_stream(pq0 + p * 4), w[p][1] = ldg_stream(pq1 + p * 4);
        dq[0] = __ldg(ps0), dq[1] = __ldg(ps1);
        pq0 += 4 * P, pq1 += 4 * P, ps0 += 1, ps1 += 1;
        if (++pf_bat == nbat) pf_bat = 0, pf_tile += gridDim.x;
    };
    uint4 wA[P][2], wB[P][2];
    uint2 dA[2], dB[2];
    if (steps > 0) load_step(wA, dA);

    pdl_wait();
    {   // source word w of (token, block): byte offset 4w = {0,16,8,24}[c] + 4*slot  ->  c = {0,2,1,3}[w >> 1], slot = w & 1
        unsigned *sb = reinterpret_cast<unsigned *>(s_b);
        const int words = (int)nb * 8;
#pragma unroll
        for (int t = 0; t < T; ++t) {
            const unsigned *src = reinterpret_cast<const unsigned *>(aq + (long)t * a_stride);
            for (int i = threadIdx.x; i < words; i += kWarps * 32) {
                const int b = i >> 3, w = i & 7, pr = w >> 1;
                const int cc = ((pr & 1) << 1) | (pr >> 1);
                sb[((b * T + t) * 4 + cc) * 2 + (w & 1)] = src[i];
            }
        }
        float *sd = reinterpret_cast<float *>(s_dy);
        for (int i = threadIdx.x; i < (int)nb * 8; i += kWarps * 32) {
            const int b = i >> 3, t = i & 7;
            sd[i] = t < T ? __int_as_float(ads[(long)t * (a_stride / QK) + b].x) : 0.f;
        }
    }
    __syncthreads();

    const int tb = g < T ? g : T - 1;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    long cu_tile = blockIdx.x;
    int cu_bat = 0, parity = 0;
    auto compute_step = [&](const uint4 (&w)[P][2], const uint2 (&dq)[2]) {
        const unsigned h0[2] = {dq[0].x, dq[0].y}, h1[2] = {dq[1].x, dq[1].y};
        const int bq = b_lo + cu_bat * 2 * P;
        const uint2 *pb = s_b + ((long)bq * T + tb) * 4 + c;
        const float2 *pd = s_dy + (long)bq * 4 + c;
#pragma unroll
        for (int p = 0; p < P; ++p) {
            unsigned f0[2][2], f1[2][2];   // [block of the pair][slot] for rows g and g+8
#pragma unroll
            for (int r = 0; r < 2; ++r) {
                const uint4 x = w[p][r];
                const unsigned sx = c < 2 ? x.z : x.x, sy = c < 2 ? x.w : x.y;
                const unsigned rx = __shfl_xor_sync(0xffffffffu, sx, 2), ry = __shfl_xor_sync(0xffffffffu, sy, 2);
                unsigned (&f)[2][2] = r ? f1 : f0;
                f[0][0] = c < 2 ? x.x : rx, f[0][1] = c < 2 ? x.y : ry;
                f[1][0] = c < 2 ? rx : x.z, f[1][1] = c < 2 ? ry : x.w;
            }
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int jb = 2 * p + j;      // block within the batch
                const uint2 bf = pb[jb * T * 4];
                const float2 dy = pd[jb * 4];
                int d[4];
                mma_s8s8_16832(d, f0[j][0], f1[j][0], f0[j][1], f1[j][1], bf.x, bf.y);
                const float dw0 = __half2float(__ushort_as_half((unsigned short)((jb & 1) ? (h0[jb >> 1] >> 16) : (h0[jb >> 1] & 0xffffu))));
                const float dw1 = __half2float(__ushort_as_half((unsigned short)((jb & 1) ? (h1[jb >> 1] >> 16) : (h1[jb >> 1] & 0xffffu))));
                acc[0] = fmaf(dw0 * dy.x, (float)d[0], acc[0]);
                acc[1] = fmaf(dw0 * dy.y, (float)d[1], acc[1]);
                acc[2] = fmaf(dw1 * dy.x, (float)d[2], acc[2]);
                acc[3] = fmaf(dw1 * dy.y, (float)d[3], acc[3]);
            }
        }
        if (++cu_bat == nbat) {
            s_red[parity][warp][lane] = make_float4(acc[0], acc[1], acc[2], acc[3]);
            __syncthreads();
            if (threadIdx.x < 128) {
                const float *p = reinterpret_cast<const float *>(&s_red[parity][0][0]) + threadIdx.x;
                float v = 0.f;
#pragma unroll
                for (int w8 = 0; w8 < kWarps; ++w8) v += p[w8 * 128];
                const int ln = threadIdx.x >> 2, i = threadIdx.x & 3;
                const long r = cu_tile * 16 + (ln >> 2) + 8 * (i >> 1);
                const int tok = 2 * (ln & 3) + (i & 1);
                if (tok < T && r < nrows) out[(long)tok * ldo + r] = v;
            }
            parity ^= 1;
            acc[0] = acc[1] = acc[2] = acc[3] = 0.f;
            cu_bat = 0, cu_tile += gridDim.x;
        }
    };
    for (long s = 0; s < steps; s += 2) {
        if (s + 1 < steps) load_step(wB, dB);
        compute_step(wA, dA);
        if (s + 1 >= steps) break;
        if (s + 2 < steps) load_step(wA, dA);
        compute_step(wB, dB);
    }
}

// ---- Q8_0 ---------------------------------------------------------------------------------------
// Unit of work = 16-byte half block (16 int8): lane-consecutive pieces are contiguous in HBM and in smem.
template <int T, int R>
__global__ void __launch_bounds__(kWarps * 32) gemv_q8_0_kernel(const uint4 *__restrict__ qs, const __half *__restrict__ sc,
                                                                long nb, long nbc, long row0, long nrows,
                                                                const int8_t *__restrict__ aq, const int2 *__restrict__ ads,
                                                                long a_stride, float *__restrict__ out, long ldo,
                                                                const tb200_gemv_item *__restrict__ items) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint4 *s_q = reinterpret_cast<uint4 *>(smem);             // [T][2*nbc]
    int2 *s_ds = reinterpret_cast<int2 *>(s_q + T * nbc * 2);  // [T][nbc]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long r_base = ((long)blockIdx.x * kWarps + warp) * R;
    const bool active = r_base < nrows;
    pdl_launch_dependents();
    pdl_wait();
    // grouped mode (mixmul): this CTA column serves one work item = one expert's rows x up to T gathered tokens
    __shared__ int s_arow[T];
    long my_row0 = row0;
    int ntok = T, orow[T];
#pragma unroll
    for (int t = 0; t < T; ++t) orow[t] = t;
    if (items) {
        const tb200_gemv_item &it = items[blockIdx.y];
        ntok = it.ntok;
        if (ntok == 0) return;   // CTA-uniform
        my_row0 = it.row0;
#pragma unroll
        for (int t = 0; t < T; ++t) orow[t] = it.out_row[min(t, ntok - 1)];
        if (threadIdx.x < T) s_arow[threadIdx.x] = it.a_row[min((int)threadIdx.x, ntok - 1)];
    } else if (threadIdx.x < T) s_arow[threadIdx.x] = threadIdx.x;
    __syncthreads();
    float acc[R][T];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int t = 0; t < T; ++t) acc[r][t] = 0.f;
    const uint4 *wrow[R];
    const __half *srow[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const long row = my_row0 + min(r_base + r, nrows - 1);
        wrow[r] = qs + row * nb * 2;
        srow[r] = sc + row * nb;
    }
    for (long b0 = 0; b0 < nb; b0 += nbc) {
        const long cnt = min(nbc, nb - b0);
        if (b0) __syncthreads();
#pragma unroll
        for (int t = 0; t < T; ++t) {
            const uint4 *src = reinterpret_cast<const uint4 *>(aq + (long)s_arow[t] * a_stride + b0 * 32);
            const int2 *dsrc = ads + (long)s_arow[t] * (a_stride / QK) + b0;
            for (int g = threadIdx.x; g < (int)cnt * 2; g += kWarps * 32) s_q[t * (int)nbc * 2 + g] = src[g];
            for (int g = threadIdx.x; g < (int)cnt; g += kWarps * 32) s_ds[t * (int)nbc + g] = dsrc[g];
        }
        __syncthreads();
        if (!active) continue;
        for (long g = lane; g < cnt * 2; g += 32) {
            uint4 w[R];
            float d[R];
#pragma unroll
            for (int r = 0; r < R; ++r) w[r] = ldg_stream(wrow[r] + b0 * 2 + g);
#pragma unroll
            for (int r = 0; r < R; ++r) d[r] = __half2float(srow[r][b0 + (g >> 1)]);
#pragma unroll
            for (int t = 0; t < T; ++t) {
                const uint4 a = s_q[t * nbc * 2 + g];
                const float dy = __int_as_float(s_ds[t * nbc + (g >> 1)].x);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    int s = dp4a_ss((int)w[r].x, (int)a.x, 0);
                    s = dp4a_ss((int)w[r].y, (int)a.y, s);
                    s = dp4a_ss((int)w[r].z, (int)a.z, s);
                    s = dp4a_ss((int)w[r].w, (int)a.w, s);
                    acc[r][t] = fmaf(d[r] * dy, (float)s, acc[r][t]);
                }
            }
        }
    }
    if (!active) return;
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int t = 0; t < T; ++t) {
            const float v = warp_sum(acc[r][t]);
            if (lane == 0 && r_base + r < nrows && t < ntok) out[(long)orow[t] * ldo + r_base + r] = v;
        }
}

// ---- G8: the generic resident form of Q4_1 / Q5_0 / Q5_K / IQ4_XS (tb200_common.cuh) -------------------------------------
// int8 quants + fp16 {a, b} per 32-block, value = a*q + b, so a block contributes a*d_y*sum(q*q_y) + b*d_y*sum(q_y): the
// integer block dot of the reference's kernels (ggml_vec_dot_q5_0_q8_0 ggml-quants.c:4498, _q4_1_q8_1 :4211 with its
// m*s term, _q5_K_q8_K :7400, _iq4_xs_q8_K :11799) with the sub-block scale folded into a. Same walk as the Q8_0 kernel.
template <int T, int R>
__global__ void __launch_bounds__(kWarps * 32) gemv_g8_kernel(const uint4 *__restrict__ qs, const uint32_t *__restrict__ sc,
                                                              long nb, long nbc, long row0, long nrows,
                                                              const int8_t *__restrict__ aq, const int2 *__restrict__ ads,
                                                              long a_stride, float *__restrict__ out, long ldo) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint4 *s_q = reinterpret_cast<uint4 *>(smem);             // [T][2*nbc]
    int2 *s_ds = reinterpret_cast<int2 *>(s_q + T * nbc * 2);  // [T][nbc]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long r_base = ((long)blockIdx.x * kWarps + warp) * R;
    const bool active = r_base < nrows;
    pdl_launch_dependents();
    pdl_wait();
    float acc[R][T];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int t = 0; t < T; ++t) acc[r][t] = 0.f;
    const uint4 *wrow[R];
    const uint32_t *srow[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const long row = row0 + min(r_base + r, nrows - 1);
        wrow[r] = qs + row * nb * 2;
        srow[r] = sc + row * nb;
    }
    for (long b0 = 0; b0 < nb; b0 += nbc) {
        const long cnt = min(nbc, nb - b0);
        if (b0) __syncthreads();
#pragma unroll
        for (int t = 0; t < T; ++t) {
            const uint4 *src = reinterpret_cast<const uint4 *>(aq + (long)t * a_stride + b0 * 32);
            const int2 *dsrc = ads + (long)t * (a_stride / QK) + b0;
            for (int g = threadIdx.x; g < (int)cnt * 2; g += kWarps * 32) s_q[t * (int)nbc * 2 + g] = src[g];
            for (int g = threadIdx.x; g < (int)cnt; g += kWarps * 32) s_ds[t * (int)nbc + g] = dsrc[g];
        }
        __syncthreads();
        if (!active) continue;
        for (long g = lane; g < cnt * 2; g += 32) {
            uint4 w[R];
            float a[R], b[R];
#pragma unroll
            for (int r = 0; r < R; ++r) w[r] = ldg_stream(wrow[r] + b0 * 2 + g);
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const uint32_t ab = __ldg(srow[r] + b0 + (g >> 1));
                a[r] = __half2float(__ushort_as_half((unsigned short)(ab & 0xffffu)));
                b[r] = (g & 1) ? 0.f : __half2float(__ushort_as_half((unsigned short)(ab >> 16)));   // the b*sum(q_y) term once per block
            }
#pragma unroll
            for (int t = 0; t < T; ++t) {
                const uint4 y = s_q[t * nbc * 2 + g];
                const int2 ds = s_ds[t * nbc + (g >> 1)];
                const float dy = __int_as_float(ds.x);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    int s = dp4a_ss((int)w[r].x, (int)y.x, 0);
                    s = dp4a_ss((int)w[r].y, (int)y.y, s);
                    s = dp4a_ss((int)w[r].z, (int)y.z, s);
                    s = dp4a_ss((int)w[r].w, (int)y.w, s);
                    acc[r][t] = fmaf(a[r] * dy, (float)s, acc[r][t]);
                    acc[r][t] = fmaf(b[r] * dy, (float)ds.y, acc[r][t]);
                }
            }
        }
    }
    if (!active) return;
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int t = 0; t < T; ++t) {
            const float v = warp_sum(acc[r][t]);
            if (lane == 0 && r_base + r < nrows) out[(long)t * ldo + r_base + r] = v;
        }
}

// ---- Q4_K ---------------------------------------------------------------------------------------
// Super-block = 128 B of nibbles = 8 pieces of 16 B; piece p = 2c+h holds, for 64-element chunk c,
// low nibbles = elements 64c+16h+[0,16) (sub-block 2c) and high nibbles = 64c+32+16h+[0,16) (sub-block 2c+1)
// (dequantize_row_q4_K, ggml-quants.c:2548-2570).
// A lane owns one whole 64-element chunk (pieces 2c and 2c+1 = 32 contiguous bytes), so 4 lanes share a super-block and
// a warp step covers 8 super-blocks = 1 KB of one row: the low nibbles of the two pieces meet activations 64c+[0,32)
// (sub-block 2c), the high nibbles 64c+[32,64) (sub-block 2c+1) — the activation row in its natural order. All 16
// six-bit scales/mins of the super-block are decoded with ~10 word-wide logic ops (get_scale_min_k4, :1891-1899, on
// four bytes at a time) and the lane picks its two pairs by a fixed shift. The first version gave every lane one
// piece and decoded two scale pairs per 16 bytes with byte-wise code: ~80 instructions per 16 B of weights, which
// capped the kernel at 3.1 TB/s (issue-bound); this form needs ~37.
// Shared memory: activations [T][super-block][chunk] as 4 x 16 B + 16 B of padding (80-byte stride: the 64-byte
// lane stride of the natural layout would be a 4-way bank conflict), then {d, sum} per 32-sub-block.
__device__ __forceinline__ void q4k_scale_min(const uint4 hdr, int j, int &sc, int &m) {   // get_scale_min_k4, :1891-1899 (blockdot checker kernel)
    // hdr.x = {d, dmin}; scales[12] = bytes 0..11 of (hdr.y, hdr.z, hdr.w)
    const uint32_t s0 = hdr.y, s1 = hdr.z, s2 = hdr.w;
    auto byte = [&](int i) -> uint32_t { const uint32_t wv = i < 4 ? s0 : (i < 8 ? s1 : s2); return (wv >> (8 * (i & 3))) & 0xff; };
    if (j < 4) { sc = byte(j) & 63; m = byte(j + 4) & 63; }
    else { sc = (byte(j + 4) & 0xF) | ((byte(j - 4) >> 6) << 4); m = (byte(j + 4) >> 4) | ((byte(j) >> 6) << 4); }
}

constexpr int kQ4KSmemPerSb = 4 * 80 + 64;   // bytes of staged activations per super-block and token

// Same launch structure as the Q4_0 kernel (multi-matrix grid, first weight batch before pdl_wait, fused all-gather
// stores); A.nb / A.nbc count super-blocks here.
template <int T, int R>
__global__ void __launch_bounds__(kWarps * 32, T * R <= 2 ? 4 : ((R == 2 || T <= 4) ? 2 : 1)) gemv_q4_K_kernel(const __grid_constant__ GemvQ4Args A) {
    constexpr int U = T * R == 1 ? 2 : 1;   // super-blocks per 4-lane group per row in flight
    extern __shared__ __align__(16) uint8_t smem[];
    const long nsb = A.nb, nsbc = A.nbc;
    uint4 *s_q = reinterpret_cast<uint4 *>(smem);                       // [T][nsbc][4 chunks][5]
    int2 *s_ds = reinterpret_cast<int2 *>(s_q + T * nsbc * 20);         // [T][nsbc*8]
    __shared__ int s_arow[T];

    pdl_launch_dependents();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int mat = 0;
#pragma unroll
    for (int m = 1; m < kMaxMats; ++m)
        if (m < A.nmat && (int)blockIdx.x >= A.cta_begin[m]) mat = m;
    const long nrows = A.nrows[mat];
    const long r_base = ((long)((int)blockIdx.x - A.cta_begin[mat]) * kWarps + warp) * R;
    const bool active = r_base < nrows;
    const uint4 *qs = (const uint4 *)A.qs[mat];
    const uint4 *sc = (const uint4 *)A.sc[mat];
    const int c = lane & 3, grp = lane >> 2;
    const int shift = 16 * (c & 1);           // this lane's scale pair inside the decoded 4-byte word

    uint4 w0[U][R], w1[U][R], hdr[U][R];
    long rowv[R];
    auto set_rows = [&](long my_row0) {
#pragma unroll
        for (int r = 0; r < R; ++r) rowv[r] = my_row0 + min(r_base + r, nrows - 1);
    };
    auto load_batch = [&](long sbase, long end) {   // super-blocks sbase + grp + 8u (row-relative), limited to < end
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long sb = sbase + grp + 8 * u;
            if (sb < end) {
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const uint4 *pw = qs + (rowv[r] * nsb + sb) * 8 + 2 * c;
                    w0[u][r] = ldg_stream(pw);
                    w1[u][r] = ldg_stream(pw + 1);
                    hdr[u][r] = __ldg(sc + rowv[r] * nsb + sb);
                }
            }
        }
    };

    int ntok = T, orow[T];
#pragma unroll
    for (int t = 0; t < T; ++t) orow[t] = t;
    if (!A.items) {
        set_rows(A.row0[mat]);
        if (active) load_batch(0, min(nsbc, nsb));
    }
    pdl_wait();
    if (A.items) {
        const tb200_gemv_item &it = A.items[blockIdx.y];
        ntok = it.ntok;
        if (ntok == 0) return;   // CTA-uniform
        set_rows(it.row0);
#pragma unroll
        for (int t = 0; t < T; ++t) orow[t] = it.out_row[min(t, ntok - 1)];
        if (threadIdx.x < T) s_arow[threadIdx.x] = it.a_row[min((int)threadIdx.x, ntok - 1)];
        if (active) load_batch(0, min(nsbc, nsb));
    } else if (threadIdx.x < T) s_arow[threadIdx.x] = threadIdx.x;
    __syncthreads();

    float acc[R][T];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int t = 0; t < T; ++t) acc[r][t] = 0.f;

    for (long sb0 = 0; sb0 < nsb; sb0 += nsbc) {
        const long cnt = min(nsbc, nsb - sb0);
        if (sb0) __syncthreads();
        const int nsbc_i = (int)nsbc;
#pragma unroll
        for (int t = 0; t < T; ++t) {
            const uint4 *src = reinterpret_cast<const uint4 *>(A.aq + (long)s_arow[t] * A.a_stride + sb0 * QK_K);
            const int2 *dsrc = A.ads + (long)s_arow[t] * (A.a_stride / QK) + sb0 * 8;
            for (int g = threadIdx.x; g < (int)cnt * 16; g += kWarps * 32)   // g = 16-byte piece; 4 pieces per chunk, chunk stride 5
                s_q[t * nsbc_i * 20 + (g >> 2) * 5 + (g & 3)] = src[g];
            for (int g = threadIdx.x; g < (int)cnt * 8; g += kWarps * 32) s_ds[t * nsbc_i * 8 + g] = dsrc[g];
        }
        __syncthreads();
        if (!active) continue;
        for (long sbb = 0; sbb < cnt; sbb += 8 * U) {
            if (sb0 | sbb) load_batch(sb0 + sbb, sb0 + cnt);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int sb = (int)sbb + grp + 8 * u;   // chunk-relative
                if (sb < (int)cnt) {
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        const uint4 h = hdr[u][r];
                        const __half2 dd = *reinterpret_cast<const __half2 *>(&h.x);
                        const float d = __low2float(dd), dmin = __high2float(dd);
                        // sub-blocks 0..3: 6 low bits of scales[0..3] / scales[4..7]; 4..7: nibbles of scales[8..11] + the 2 high bits
                        const uint32_t scw = c < 2 ? (h.y & 0x3f3f3f3fu) : ((h.w & 0x0f0f0f0fu) | ((h.y >> 2) & 0x30303030u));
                        const uint32_t mw = c < 2 ? (h.z & 0x3f3f3f3fu) : (((h.w >> 4) & 0x0f0f0f0fu) | ((h.z >> 2) & 0x30303030u));
                        const int sc_lo = (scw >> shift) & 0xff, sc_hi = (scw >> (shift + 8)) & 0xff;
                        const int m_lo = (mw >> shift) & 0xff, m_hi = (mw >> (shift + 8)) & 0xff;
                        const uint4 a = w0[u][r], b = w1[u][r];
                        const unsigned al0 = a.x & 0x0f0f0f0fu, al1 = a.y & 0x0f0f0f0fu, al2 = a.z & 0x0f0f0f0fu, al3 = a.w & 0x0f0f0f0fu;
                        const unsigned bl0 = b.x & 0x0f0f0f0fu, bl1 = b.y & 0x0f0f0f0fu, bl2 = b.z & 0x0f0f0f0fu, bl3 = b.w & 0x0f0f0f0fu;
                        const unsigned ah0 = a.x & 0xf0f0f0f0u, ah1 = a.y & 0xf0f0f0f0u, ah2 = a.z & 0xf0f0f0f0u, ah3 = a.w & 0xf0f0f0f0u;   // 16*q
                        const unsigned bh0 = b.x & 0xf0f0f0f0u, bh1 = b.y & 0xf0f0f0f0u, bh2 = b.z & 0xf0f0f0f0u, bh3 = b.w & 0xf0f0f0f0u;
#pragma unroll
                        for (int t = 0; t < T; ++t) {
                            const uint4 *ap = s_q + ((t * nsbc_i + sb) * 4 + c) * 5;
                            const uint4 y0 = ap[0], y1 = ap[1], y2 = ap[2], y3 = ap[3];
                            int sl = dp4a_us(al0, (int)y0.x, 0);
                            sl = dp4a_us(al1, (int)y0.y, sl), sl = dp4a_us(al2, (int)y0.z, sl), sl = dp4a_us(al3, (int)y0.w, sl);
                            sl = dp4a_us(bl0, (int)y1.x, sl), sl = dp4a_us(bl1, (int)y1.y, sl), sl = dp4a_us(bl2, (int)y1.z, sl), sl = dp4a_us(bl3, (int)y1.w, sl);
                            int sh = dp4a_us(ah0, (int)y2.x, 0);
                            sh = dp4a_us(ah1, (int)y2.y, sh), sh = dp4a_us(ah2, (int)y2.z, sh), sh = dp4a_us(ah3, (int)y2.w, sh);
                            sh = dp4a_us(bh0, (int)y3.x, sh), sh = dp4a_us(bh1, (int)y3.y, sh), sh = dp4a_us(bh2, (int)y3.z, sh), sh = dp4a_us(bh3, (int)y3.w, sh);
                            const int4 dsv = *reinterpret_cast<const int4 *>(s_ds + (t * nsbc_i + sb) * 8 + 2 * c);   // {d, sum} of sub-blocks 2c, 2c+1
                            const float dy = __int_as_float(dsv.x);
                            const int isum = sc_lo * sl + sc_hi * (sh >> 4);   // sh is an exact multiple of 16
                            const int msum = m_lo * dsv.y + m_hi * dsv.w;
                            acc[r][t] = fmaf(d * dy, (float)isum, acc[r][t]);
                            acc[r][t] = fmaf(-(dmin * dy), (float)msum, acc[r][t]);
                        }
                    }
                }
            }
        }
    }
    store_rows<T, R>(A, mat, acc, orow, ntok, r_base, nrows, lane, active);
}

// ---- Q6_K ---------------------------------------------------------------------------------------
// Resident layout (tb200_repack.cu): per super-block 4 units of 48 B = three 16-byte pieces; unit u = 2*half + r covers
// the 64 elements 128*half + 16*r + 32*j + [0,16), j = 0..3, one int8 scale per j: scales[8*half + r + 2j], stored as
// byte 4u + j (dequantize_row_q6_K, ggml-quants.c:2970-3000; dot: ggml_vec_dot_q6_K_q8_K scalar branch :8642-8682,
// served by iqk DequantizerQ6K, llamafile/iqk_mul_mat.inc:570-600). Piece j < 3 holds q_j << 2 with two bits of q_3 in
// the low end of every byte, so (piece & 0xfc) ^ 0x80 is 4*(q_j - 32) as int8 with no nibble/high-bit merge: 9 logic
// ops per 16 weights instead of 28 for the file's ql/qh split. A lane owns one unit (3 x 16-byte loads, 48 contiguous
// bytes), 4 lanes share a super-block, a warp step covers 8 super-blocks = 1.5 KB of one row. dp4a sums are 4x the
// reference's integers (exact; the 1/4 is folded into d).
// Shared memory: activations [T][super-block][j][u] as 16-byte pieces, 20 pieces per super-block (4 of padding) so the
// 8 lanes of a quarter-warp (2 super-blocks x 4 units) hit 8 distinct bank groups; then the Q8_K scale per sub-block.
constexpr int kQ6KSmemPerSb = 20 * 16 + 64;

// Occupancy over per-thread depth: with <= 64 registers four CTAs fit an SM, which measured 68 us on the 128256-row matrix
// against 83 us for the two-CTA version that kept two super-blocks per row in flight (and 91 us for five CTAs at 48
// registers, which spills).
template <int T, int R>
__global__ void __launch_bounds__(kWarps * 32, T * R <= 2 ? 4 : (T == 3 && R == 1 ? 3 : 2)) gemv_q6_K_kernel(const __grid_constant__ GemvQ4Args A) {
    constexpr int U = T * R == 1 ? 2 : 1;   // super-blocks per 4-lane group per row in flight
    extern __shared__ __align__(16) uint8_t smem[];
    const long nsb = A.nb, nsbc = A.nbc;
    uint4 *s_q = reinterpret_cast<uint4 *>(smem);                       // [T][nsbc][20]
    int2 *s_ds = reinterpret_cast<int2 *>(s_q + T * nsbc * 20);         // [T][nsbc*8]
    __shared__ int s_arow[T];

    pdl_launch_dependents();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int mat = 0;
#pragma unroll
    for (int m = 1; m < kMaxMats; ++m)
        if (m < A.nmat && (int)blockIdx.x >= A.cta_begin[m]) mat = m;
    const long nrows = A.nrows[mat];
    const long r_base = ((long)((int)blockIdx.x - A.cta_begin[mat]) * kWarps + warp) * R;
    const bool active = r_base < nrows;
    const uint4 *qs = (const uint4 *)A.qs[mat];
    const uint32_t *sc = (const uint32_t *)A.sc[mat];   // word u of a super-block = the 4 int8 scales of unit u
    const __half *dp = (const __half *)A.d6[mat];
    const int u = lane & 3, grp = lane >> 2;

    uint4 wa[U][R], wb[U][R], wh[U][R];   // pieces j = 0, 1, 2: byte = q_j << 2 | two bits of q_3 (tb200_repack.cu)
    uint32_t hdr[U][R];
    __half dw[U][R];
    long rowv[R];
    auto set_rows = [&](long my_row0) {
#pragma unroll
        for (int r = 0; r < R; ++r) rowv[r] = my_row0 + min(r_base + r, nrows - 1);
    };
    auto load_batch = [&](long sbase, long end) {   // super-blocks sbase + grp + 8u' (row-relative), limited to < end
#pragma unroll
        for (int uu = 0; uu < U; ++uu) {
            const long sb = sbase + grp + 8 * uu;
            if (sb < end) {
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const uint4 *pw = qs + (rowv[r] * nsb + sb) * 12 + 3 * u;
                    wa[uu][r] = ldg_stream(pw);
                    wb[uu][r] = ldg_stream(pw + 1);
                    wh[uu][r] = ldg_stream(pw + 2);
                    hdr[uu][r] = __ldg(sc + (rowv[r] * nsb + sb) * 4 + u);
                    dw[uu][r] = dp[rowv[r] * nsb + sb];
                }
            }
        }
    };

    int ntok = T, orow[T];
#pragma unroll
    for (int t = 0; t < T; ++t) orow[t] = t;
    if (!A.items) {
        set_rows(A.row0[mat]);
        if (active) load_batch(0, min(nsbc, nsb));
    }
    pdl_wait();
    if (A.items) {
        const tb200_gemv_item &it = A.items[blockIdx.y];
        ntok = it.ntok;
        if (ntok == 0) return;   // CTA-uniform
        set_rows(it.row0);
#pragma unroll
        for (int t = 0; t < T; ++t) orow[t] = it.out_row[min(t, ntok - 1)];
        if (threadIdx.x < T) s_arow[threadIdx.x] = it.a_row[min((int)threadIdx.x, ntok - 1)];
        if (active) load_batch(0, min(nsbc, nsb));
    } else if (threadIdx.x < T) s_arow[threadIdx.x] = threadIdx.x;
    __syncthreads();

    float acc[R][T];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int t = 0; t < T; ++t) acc[r][t] = 0.f;

    for (long sb0 = 0; sb0 < nsb; sb0 += nsbc) {
        const long cnt = min(nsbc, nsb - sb0);
        if (sb0) __syncthreads();
        const int nsbc_i = (int)nsbc;
#pragma unroll
        for (int t = 0; t < T; ++t) {
            const uint4 *src = reinterpret_cast<const uint4 *>(A.aq + (long)s_arow[t] * A.a_stride + sb0 * QK_K);
            const int2 *dsrc = A.ads + (long)s_arow[t] * (A.a_stride / QK) + sb0 * 8;
            for (int g = threadIdx.x; g < (int)cnt * 16; g += kWarps * 32) {
                // natural piece q = 8*half + 2j + r of its super-block -> slot 4j + (2*half + r)
                const int q = g & 15, pos = ((q & 7) >> 1) * 4 + ((q >> 3) * 2 + (q & 1));
                s_q[(t * nsbc_i + (g >> 4)) * 20 + pos] = src[g];
            }
            for (int g = threadIdx.x; g < (int)cnt * 8; g += kWarps * 32) s_ds[t * nsbc_i * 8 + g] = dsrc[g];
        }
        __syncthreads();
        if (!active) continue;
        for (long sbb = 0; sbb < cnt; sbb += 8 * U) {
            if (sb0 | sbb) load_batch(sb0 + sbb, sb0 + cnt);
#pragma unroll
            for (int uu = 0; uu < U; ++uu) {
                const int sb = (int)sbb + grp + 8 * uu;   // chunk-relative
                if (sb < (int)cnt) {
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        const uint32_t h = hdr[uu][r];
                        const float d = 0.25f * __half2float(dw[uu][r]);   // the int8 operands below are 4*(q - 32)
                        const int sc0 = (int)(int8_t)h, sc1 = (int)(int8_t)(h >> 8), sc2 = (int)(int8_t)(h >> 16), sc3 = (int)h >> 24;
                        const uint32_t a[4] = {wa[uu][r].x, wa[uu][r].y, wa[uu][r].z, wa[uu][r].w};
                        const uint32_t b[4] = {wb[uu][r].x, wb[uu][r].y, wb[uu][r].z, wb[uu][r].w};
                        const uint32_t hh[4] = {wh[uu][r].x, wh[uu][r].y, wh[uu][r].z, wh[uu][r].w};
                        uint32_t q0[4], q1[4], q2[4], q3[4];
#pragma unroll
                        for (int w = 0; w < 4; ++w) {   // top 6 bits of a byte ^ 0x80 = int8 4*(q - 32); q_3 from the three 2-bit tails
                            q0[w] = (a[w] & 0xfcfcfcfcu) ^ 0x80808080u;
                            q1[w] = (b[w] & 0xfcfcfcfcu) ^ 0x80808080u;
                            q2[w] = (hh[w] & 0xfcfcfcfcu) ^ 0x80808080u;
                            q3[w] = ((((a[w] << 6) & 0xc0c0c0c0u) ^ 0x80808080u) | ((b[w] << 4) & 0x30303030u)) | ((hh[w] << 2) & 0x0c0c0c0cu);
                        }
#pragma unroll
                        for (int t = 0; t < T; ++t) {
                            const uint4 *ap = s_q + (t * nsbc_i + sb) * 20 + u;
                            const uint4 y0 = ap[0], y1 = ap[4], y2 = ap[8], y3 = ap[12];
                            int d0 = dp4a_ss((int)q0[0], (int)y0.x, 0), d1 = dp4a_ss((int)q1[0], (int)y1.x, 0);
                            int d2 = dp4a_ss((int)q2[0], (int)y2.x, 0), d3 = dp4a_ss((int)q3[0], (int)y3.x, 0);
                            d0 = dp4a_ss((int)q0[1], (int)y0.y, d0), d1 = dp4a_ss((int)q1[1], (int)y1.y, d1);
                            d2 = dp4a_ss((int)q2[1], (int)y2.y, d2), d3 = dp4a_ss((int)q3[1], (int)y3.y, d3);
                            d0 = dp4a_ss((int)q0[2], (int)y0.z, d0), d1 = dp4a_ss((int)q1[2], (int)y1.z, d1);
                            d2 = dp4a_ss((int)q2[2], (int)y2.z, d2), d3 = dp4a_ss((int)q3[2], (int)y3.z, d3);
                            d0 = dp4a_ss((int)q0[3], (int)y0.w, d0), d1 = dp4a_ss((int)q1[3], (int)y1.w, d1);
                            d2 = dp4a_ss((int)q2[3], (int)y2.w, d2), d3 = dp4a_ss((int)q3[3], (int)y3.w, d3);
                            const float dy = __int_as_float(s_ds[(t * nsbc_i + sb) * 8].x);   // Q8_K: one d per super-block
                            const int isum = sc0 * d0 + sc1 * d1 + sc2 * d2 + sc3 * d3;
                            acc[r][t] = fmaf(d * dy, (float)isum, acc[r][t]);
                        }
                    }
                }
            }
        }
    }
    store_rows<T, R>(A, mat, acc, orow, ntok, r_base, nrows, lane, active);
}

// ---- F16 ----------------------------------------------------------------------------------------
template <int T, int R, bool kActF32>
__global__ void __launch_bounds__(kWarps * 32) gemv_f16_kernel(const __half *__restrict__ wt, long k, long kc, long row0, long nrows,
                                                               const void *__restrict__ act, long a_stride,
                                                               float *__restrict__ out, long ldo,
                                                                const tb200_gemv_item *__restrict__ items) {
    extern __shared__ __align__(16) uint8_t smem[];
    float *s_a = reinterpret_cast<float *>(smem);   // [T][kc] widened to fp32 once per chunk
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long r_base = ((long)blockIdx.x * kWarps + warp) * R;
    const bool active = r_base < nrows;
    pdl_launch_dependents();
    pdl_wait();
    // grouped mode (mixmul): this CTA column serves one work item = one expert's rows x up to T gathered tokens
    __shared__ int s_arow[T];
    long my_row0 = row0;
    int ntok = T, orow[T];
#pragma unroll
    for (int t = 0; t < T; ++t) orow[t] = t;
    if (items) {
        const tb200_gemv_item &it = items[blockIdx.y];
        ntok = it.ntok;
        if (ntok == 0) return;   // CTA-uniform
        my_row0 = it.row0;
#pragma unroll
        for (int t = 0; t < T; ++t) orow[t] = it.out_row[min(t, ntok - 1)];
        if (threadIdx.x < T) s_arow[threadIdx.x] = it.a_row[min((int)threadIdx.x, ntok - 1)];
    } else if (threadIdx.x < T) s_arow[threadIdx.x] = threadIdx.x;
    __syncthreads();
    float acc[R][T];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int t = 0; t < T; ++t) acc[r][t] = 0.f;
    const bool vec = (k & 7) == 0 && (kc & 7) == 0 && ((uintptr_t)wt & 15) == 0;
    for (long k0 = 0; k0 < k; k0 += kc) {
        const long cnt = min(kc, k - k0);
        if (k0) __syncthreads();
#pragma unroll
        for (int t = 0; t < T; ++t)
            for (int j = threadIdx.x; j < (int)cnt; j += kWarps * 32)
                s_a[t * kc + j] = kActF32 ? reinterpret_cast<const float *>(act)[(long)s_arow[t] * a_stride + k0 + j]
                                          : __half2float(reinterpret_cast<const __half *>(act)[(long)s_arow[t] * a_stride + k0 + j]);
        __syncthreads();
        if (!active) continue;
        if (vec) {
            for (long g = lane; g < cnt / 8; g += 32) {
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const long row = my_row0 + min(r_base + r, nrows - 1);
                    const uint4 w = ldg_stream(reinterpret_cast<const uint4 *>(wt + row * k + k0) + g);
                    const __half2 *hp = reinterpret_cast<const __half2 *>(&w);
                    float wf[8];
#pragma unroll
                    for (int i = 0; i < 4; ++i) { const float2 f = __half22float2(hp[i]); wf[2 * i] = f.x, wf[2 * i + 1] = f.y; }
#pragma unroll
                    for (int t = 0; t < T; ++t) {
                        const float4 a0 = *reinterpret_cast<const float4 *>(s_a + t * kc + g * 8), a1 = *reinterpret_cast<const float4 *>(s_a + t * kc + g * 8 + 4);
                        float s = acc[r][t];
                        s = fmaf(wf[0], a0.x, s), s = fmaf(wf[1], a0.y, s), s = fmaf(wf[2], a0.z, s), s = fmaf(wf[3], a0.w, s);
                        s = fmaf(wf[4], a1.x, s), s = fmaf(wf[5], a1.y, s), s = fmaf(wf[6], a1.z, s), s = fmaf(wf[7], a1.w, s);
                        acc[r][t] = s;
                    }
                }
            }
        } else {
            for (long j = lane; j < cnt; j += 32)
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const long row = my_row0 + min(r_base + r, nrows - 1);
                    const float wv = __half2float(wt[row * k + k0 + j]);
#pragma unroll
                    for (int t = 0; t < T; ++t) acc[r][t] = fmaf(wv, s_a[t * kc + j], acc[r][t]);
                }
        }
    }
    if (!active) return;
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int t = 0; t < T; ++t) {
            const float v = warp_sum(acc[r][t]);
            if (lane == 0 && r_base + r < nrows && t < ntok) out[(long)orow[t] * ldo + r_base + r] = v;
        }
}

// ---- exact integer block dots (parity instrument; not a hot path) -------------------------------
__global__ void blockdot_kernel(int type, const uint8_t *__restrict__ qs, const uint8_t *__restrict__ sc, long rows, long k,
                                const int8_t *__restrict__ aq, const int2 *__restrict__ ads, int32_t *__restrict__ out) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (type == TB200_TYPE_Q6_K) {   // one int32 per super-block: sum_j scale_j * sum (q6 - 32) * q8, from the resident unit layout
        const long nsb = k / QK_K;
        if (i >= rows * nsb) return;
        const long sb = i % nsb;
        const uint8_t *q = qs + i * 192;
        const int8_t *scl = reinterpret_cast<const int8_t *>(sc) + i * 16;   // byte 4u + j
        int sum = 0;
        for (int u = 0; u < 4; ++u) {
            const int half = u >> 1, r = u & 1;
            const uint8_t *p0 = q + u * 48, *p1 = p0 + 16, *p2 = p0 + 32;
            for (int j = 0; j < 4; ++j) {
                int dot = 0;
                for (int l = 0; l < 16; ++l) {
                    const int q6 = j == 0 ? (p0[l] >> 2) : j == 1 ? (p1[l] >> 2) : j == 2 ? (p2[l] >> 2)
                                                                                 : (((p0[l] & 3) << 4) | ((p1[l] & 3) << 2) | (p2[l] & 3));
                    dot += (q6 - 32) * aq[sb * 256 + 128 * half + 16 * r + 32 * j + l];
                }
                sum += (int)scl[4 * u + j] * dot;
            }
        }
        out[i] = sum;
        return;
    }
    if (type == TB200_TYPE_Q4_K) {
        const long nsb = k / QK_K;
        if (i >= rows * nsb) return;
        const long sb = i % nsb;
        const uint8_t *q4 = qs + i * 128;
        const uint4 hdr = *reinterpret_cast<const uint4 *>(sc + i * 16);
        int sum_q = 0, sum_m = 0;
        for (int s = 0; s < 8; ++s) {
            int scs, ms;
            q4k_scale_min(hdr, s, scs, ms);
            int dot = 0;
            for (int l = 0; l < 32; ++l) {
                const int q = (s & 1) ? (q4[(s / 2) * 32 + l] >> 4) : (q4[(s / 2) * 32 + l] & 0xF);
                dot += q * aq[sb * 256 + s * 32 + l];
            }
            sum_q += scs * dot;
            sum_m += ms * ads[sb * 8 + s].y;
        }
        out[2 * i] = sum_q, out[2 * i + 1] = sum_m;
        return;
    }
    const long nb = k / QK;
    if (i >= rows * nb) return;
    const long b = i % nb;
    int s = 0;
    if (type == TB200_TYPE_Q4_0) {
        const uint8_t *q4 = qs + i * 16;
        for (int j = 0; j < 16; ++j) s += ((q4[j] & 0xF) - 8) * aq[b * 32 + j] + ((q4[j] >> 4) - 8) * aq[b * 32 + 16 + j];
    } else {
        const int8_t *q8 = reinterpret_cast<const int8_t *>(qs) + i * 32;
        for (int j = 0; j < 32; ++j) s += q8[j] * aq[b * 32 + j];
    }
    out[i] = s;
}

template <typename K>
int set_smem(K kernel, size_t bytes) {
    if (bytes > 47 * 1024) {   // static shared memory counts against the 48 KB default too
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
        if (e != cudaSuccess) { tb200_set_error("gemv: %zu bytes of shared memory: %s", bytes, cudaGetErrorString(e)); return TB200_ERR_CUDA; }
    }
    return TB200_OK;
}

// kernel arguments of the Q4_0 / Q4_K decode kernels from the launcher-level request: concatenated row tiles of up to
// kMaxMats matrices, optional peer bases for the fused all-gather
template <int R>
int fill_args(GemvQ4Args &ka, const tb200_gemv_args &a, long nb, long nbc) {
    memset(&ka, 0, sizeof(ka));
    ka.nmat = a.nmat > 0 ? a.nmat : 1;
    int ctas = 0;
    for (int m = 0; m < ka.nmat; ++m) {
        const tb200_weights *wm = m == 0 ? a.w : a.w_more[m - 1];
        const long nr = m == 0 ? a.nrows : wm->rows;
        ka.cta_begin[m] = ctas;
        ctas += (int)((nr + kWarps * R - 1) / (kWarps * R));
        ka.qs[m] = wm->qs, ka.sc[m] = wm->sc, ka.d6[m] = wm->d;
        ka.row0[m] = m == 0 ? a.row0 : 0, ka.nrows[m] = nr;
        ka.out[m] = m == 0 ? a.out : a.out_more[m - 1], ka.ldo[m] = m == 0 ? a.ldo : a.ldo_more[m - 1];
    }
    ka.cta_begin[ka.nmat] = ctas;
    ka.nb = nb, ka.nbc = nbc, ka.aq = a.aq, ka.ads = a.ads, ka.a_stride = a.a_stride, ka.items = a.items, ka.n_live = a.n_live;
    ka.world = a.world > 1 ? a.world : 1;
    ka.local_base = (const char *)a.local_base;
    for (int p = 0; p < TB200_MAX_PEERS; ++p) ka.peer_base[p] = p < ka.world ? (char *)a.peer_base[p] : nullptr;
    return ctas;
}

constexpr size_t kSmemBudget = 96 * 1024;   // activation staging per CTA; larger T*K is walked in K chunks

template <int T, int R>
int launch_TR(tb200_ctx *c, const tb200_gemv_args &a) {
    const tb200_weights *w = a.w;
    const dim3 grid((unsigned)((a.nrows + kWarps * R - 1) / (kWarps * R)), a.items ? (unsigned)a.n_items : 1u);
    int rc;
    switch (w->type) {
    case TB200_TYPE_Q4_0:
    case TB200_TYPE_Q8_0: {
        const long nb = w->k / QK;
        long nbc = (long)(kSmemBudget / ((size_t)T * 40)) / 32 * 32;
        if (nbc >= nb) nbc = nb;
        const size_t smem = (size_t)T * nbc * 40;
        if (w->type == TB200_TYPE_Q4_0) {
            GemvQ4Args ka;
            const int ctas = fill_args<R>(ka, a, nb, nbc);
            const dim3 g2((unsigned)ctas, grid.y);
            if ((rc = set_smem(gemv_q4_0_kernel<T, R>, smem))) return rc;
            TB_CUDA(tb200_launch(gemv_q4_0_kernel<T, R>, g2, dim3(kWarps * 32), smem, c->stream, ka));
        } else {
            if ((rc = set_smem(gemv_q8_0_kernel<T, R>, smem))) return rc;
            TB_CUDA(tb200_launch(gemv_q8_0_kernel<T, R>, grid, dim3(kWarps * 32), smem, c->stream, (const uint4 *)w->qs, (const __half *)w->sc, nb, nbc, a.row0, a.nrows,
                                                                          a.aq, a.ads, a.a_stride, a.out, a.ldo, a.items));
        }
        break;
    }
    case TB200_TYPE_Q4_K: {
        const long nsb = w->k / QK_K;
        long nsbc = (long)(kSmemBudget / ((size_t)T * kQ4KSmemPerSb)) / 8 * 8;
        if (nsbc >= nsb) nsbc = nsb;
        const size_t smem = (size_t)T * nsbc * kQ4KSmemPerSb;
        GemvQ4Args ka;
        const int ctas = fill_args<R>(ka, a, nsb, nsbc);
        const dim3 g2((unsigned)ctas, grid.y);
        if ((rc = set_smem(gemv_q4_K_kernel<T, R>, smem))) return rc;
        TB_CUDA(tb200_launch(gemv_q4_K_kernel<T, R>, g2, dim3(kWarps * 32), smem, c->stream, ka));
        break;
    }
    case TB200_TYPE_Q6_K: {
        const long nsb = w->k / QK_K;
        long nsbc = (long)(kSmemBudget / ((size_t)T * kQ6KSmemPerSb)) / 8 * 8;
        if (nsbc >= nsb) nsbc = nsb;
        const size_t smem = (size_t)T * nsbc * kQ6KSmemPerSb;
        GemvQ4Args ka;
        const int ctas = fill_args<R>(ka, a, nsb, nsbc);
        const dim3 g2((unsigned)ctas, grid.y);
        if ((rc = set_smem(gemv_q6_K_kernel<T, R>, smem))) return rc;
        TB_CUDA(tb200_launch(gemv_q6_K_kernel<T, R>, g2, dim3(kWarps * 32), smem, c->stream, ka));
        break;
    }
    case TB200_TYPE_Q4_1: case TB200_TYPE_Q5_0: case TB200_TYPE_Q5_K: case TB200_TYPE_IQ4_XS: {
        if (a.items) { tb200_set_error("gemv: grouped mode is not available for type %d", w->type); return TB200_ERR_UNSUPPORTED; }
        const long nb = w->k / QK;
        long nbc = (long)(kSmemBudget / ((size_t)T * 40)) / 32 * 32;
        if (nbc >= nb) nbc = nb;
        const size_t smem = (size_t)T * nbc * 40;
        if ((rc = set_smem(gemv_g8_kernel<T, R>, smem))) return rc;
        TB_CUDA(tb200_launch(gemv_g8_kernel<T, R>, grid, dim3(kWarps * 32), smem, c->stream, (const uint4 *)w->qs, (const uint32_t *)w->sc, nb, nbc, a.row0, a.nrows,
                             a.aq, a.ads, a.a_stride, a.out, a.ldo));
        break;
    }
    case TB200_TYPE_F16: {
        long kc = (long)(kSmemBudget / ((size_t)T * 4)) / 256 * 256;
        if (kc >= w->k) kc = w->k;
        const size_t smem = (size_t)T * kc * 4;
        if (a.af_is_f32) {
            if ((rc = set_smem(gemv_f16_kernel<T, R, true>, smem))) return rc;
            TB_CUDA(tb200_launch(gemv_f16_kernel<T, R, true>, grid, dim3(kWarps * 32), smem, c->stream, (const __half *)w->qs, w->k, kc, a.row0, a.nrows, a.af, a.a_stride, a.out, a.ldo, a.items));
        } else {
            if ((rc = set_smem(gemv_f16_kernel<T, R, false>, smem))) return rc;
            TB_CUDA(tb200_launch(gemv_f16_kernel<T, R, false>, grid, dim3(kWarps * 32), smem, c->stream, (const __half *)w->qs, w->k, kc, a.row0, a.nrows, a.af, a.a_stride, a.out, a.ldo, a.items));
        }
        break;
    }
    default:
        tb200_set_error("gemv: unsupported weight type %d", w->type);
        return TB200_ERR_UNSUPPORTED;
    }
    return TB200_OK;
}

bool resident_enabled() {
    static const int on = [] { const char *v = getenv("TB200_GEMV_RESIDENT"); return (v && *v == '0') ? 0 : 1; }();
    return on != 0;
}

// Q4_0 dense launches with more row pairs than resident warps: the resident-grid kernel (3 CTAs per SM)
template <int T>
int launch_resident_q4_0(tb200_ctx *c, const tb200_gemv_args &a) {
    const long nb = a.w->k / QK;
    GemvQ4Args ka;
    fill_args<2>(ka, a, nb, nb);
    long units = 0;   // cta_begin[] carries row-pair offsets for this kernel
    for (int m = 0; m < ka.nmat; ++m) { ka.cta_begin[m] = (int)units; units += (ka.nrows[m] + 1) / 2; }
    ka.cta_begin[ka.nmat] = (int)units;
    const size_t smem = (size_t)T * nb * 40;
    int rc;
    if ((rc = set_smem(gemv_q4_0_resident_kernel<T>, smem))) return rc;
    long ctas = (units + kWarps - 1) / kWarps;
    if (ctas > 3L * c->sm_count) ctas = 3L * c->sm_count;
    TB_CUDA(tb200_launch(gemv_q4_0_resident_kernel<T>, dim3((unsigned)ctas), dim3(kWarps * 32), smem, c->stream, ka));
    return TB200_OK;
}

// grouped Q4_0 launches whose item table leads with its live items (mixmul decode): resident grid, item chosen per CTA
size_t mma_q4_0_smem(int T, long k) { return (size_t)(k / QK) * (T * 32 + 64); }
bool mma_enabled() {
    static const int on = [] { const char *v = getenv("TB200_GEMV_MMA"); return (v && *v == '0') ? 0 : 1; }();
    return on != 0;
}

// dense Q4_0 launches with 2..8 tokens: integer tensor-core kernel on a resident grid
template <int T>
int launch_mma_q4_0(tb200_ctx *c, const tb200_gemv_args &a) {
    const long nb = a.w->k / QK;
    GemvQ4Args ka;
    fill_args<2>(ka, a, nb, nb);
    const size_t smem = mma_q4_0_smem(T, a.w->k);
    const long tiles = (a.nrows + 15) / 16;
    // a warp per tile needs >= 3 rounds of tiles for the resident warps to balance; smaller matrices split K over the CTA's warps
    // (a warp per tile, KS = 1, was measured on the vocabulary matrix: 107 us against 100 us split - 8016 tiles are 2.26
    // rounds of the 3552 resident warps - so every Llama shape takes the split form; KS = 1 is kept for taller matrices)
    static const int force = [] { const char *v = getenv("TB200_MMA_SPLIT"); return v && *v ? atoi(v) : -1; }();
    const bool split = force >= 0 ? force != 0 : tiles < 6L * 3 * c->sm_count * kWarps;
    const size_t per_cta = smem + 1024 + (split ? 2 * kWarps * 32 * 16 : 512);
    const int resident = per_cta * 3 <= 227 * 1024 ? 3 : (per_cta * 2 <= 227 * 1024 ? 2 : 1);
    long ctas = split ? tiles : (tiles + kWarps - 1) / kWarps;
    if (ctas > (long)resident * c->sm_count) ctas = (long)resident * c->sm_count;
    int rc;
    if (split) {
        static const bool pair = [] { const char *v = getenv("TB200_MMA_PAIR"); return v && *v == '1'; }();
        static const bool pref = [] { const char *v = getenv("TB200_MMA_PREFETCH"); return !(v && *v == '0'); }();   // N=14336 T=4 15.4 -> 14.3 us, K=14336 T=8 29.2 -> 25.4 us, vocabulary matrix unchanged
        if (pref) ka.flags |= 1;
        if (pair) {
            if ((rc = set_smem(gemm_q4_0_mma_kernel<T, kWarps, true>, smem))) return rc;
            TB_CUDA(tb200_launch(gemm_q4_0_mma_kernel<T, kWarps, true>, dim3((unsigned)ctas), dim3(kWarps * 32), smem, c->stream, ka));
            return TB200_OK;
        }
        if ((rc = set_smem(gemm_q4_0_mma_kernel<T, kWarps>, smem))) return rc;
        TB_CUDA(tb200_launch(gemm_q4_0_mma_kernel<T, kWarps>, dim3((unsigned)ctas), dim3(kWarps * 32), smem, c->stream, ka));
    } else {
        if ((rc = set_smem(gemm_q4_0_mma_kernel<T, 1>, smem))) return rc;
        TB_CUDA(tb200_launch(gemm_q4_0_mma_kernel<T, 1>, dim3((unsigned)ctas), dim3(kWarps * 32), smem, c->stream, ka));
    }
    return TB200_OK;
}

template <int T>
int launch_mma_q8_0(tb200_ctx *c, const tb200_gemv_args &a) {
    const long nb = a.w->k / QK;
    const size_t smem = (size_t)nb * (T * 32 + 32);
    const long tiles = (a.nrows + 15) / 16;
    const size_t per_cta = smem + 1024 + 2 * kWarps * 32 * 16;
    const int resident = per_cta * 3 <= 227 * 1024 ? 3 : (per_cta * 2 <= 227 * 1024 ? 2 : 1);
    long ctas = tiles;
    if (ctas > (long)resident * c->sm_count) ctas = (long)resident * c->sm_count;
    int rc;
    if ((rc = set_smem(gemm_q8_0_mma_kernel<T>, smem))) return rc;
    TB_CUDA(tb200_launch(gemm_q8_0_mma_kernel<T>, dim3((unsigned)ctas), dim3(kWarps * 32), smem, c->stream, (const uint4 *)a.w->qs, (const __half *)a.w->sc, nb,
                         a.row0, a.nrows, a.aq, a.ads, a.a_stride, a.out, a.ldo));
    return TB200_OK;
}

size_t grouped_q4_0_smem(int T, long k) { return (size_t)(k / QK) * (T * 40 + 16); }

template <int T>
int launch_grouped_q4_0(tb200_ctx *c, const tb200_gemv_args &a) {
    const long nb = a.w->k / QK;
    GemvQ4Args ka;
    fill_args<2>(ka, a, nb, nb);
    const size_t smem = grouped_q4_0_smem(T, a.w->k);
    int rc;
    if ((rc = set_smem(gemv_q4_0_grouped_kernel<T>, smem))) return rc;
    const size_t per_cta = smem + 1040 + 1024;   // + the kernel's static shared memory + the per-CTA reservation
    const int resident = per_cta * 2 <= 227 * 1024 ? 2 : 1;   // the kernel is compiled for two CTAs per SM (128 registers)
    TB_CUDA(tb200_launch(gemv_q4_0_grouped_kernel<T>, dim3((unsigned)(resident * c->sm_count)), dim3(kWarps * 32), smem, c->stream, ka));
    return TB200_OK;
}

template <int T>
int launch_T(tb200_ctx *c, const tb200_gemv_args &a) {
    // rows per warp: 2 keeps >= 8 x 16-byte weight loads in flight per lane at K = 4096 while leaving
    // enough CTAs (N/8) to cover 148 SMs; small matrices fall back to 1 row per warp.
    long total = a.nrows * (a.items ? (a.n_items > 0 ? a.n_items : 1) : 1);   // grouped mode: every item walks nrows rows
    for (int m = 1; m < a.nmat; ++m) total += a.w_more[m - 1]->rows;
    if constexpr (T >= 4)   // below 4 tokens the dp4a kernels are faster (vocabulary matrix, T = 3: 89 vs 99 us)
        if (a.w->type == TB200_TYPE_Q4_0 && !a.items && a.nmat <= 1 && a.world <= 1 && mma_enabled() && (a.w->k / QK) % 32 == 0 &&
            mma_q4_0_smem(T, a.w->k) <= 200 * 1024)
            return launch_mma_q4_0<T>(c, a);
    if constexpr (T >= 2)   // Q8_0 streams twice the bytes per mma, so the tensor kernel already wins at 2 tokens (vocabulary matrix: 89 vs 118 us)
        if (a.w->type == TB200_TYPE_Q8_0 && !a.items && a.nmat <= 1 && a.world <= 1 && mma_enabled() && (a.w->k / QK) % 32 == 0 &&
            (size_t)(a.w->k / QK) * (T * 32 + 32) <= 200 * 1024)
            return launch_mma_q8_0<T>(c, a);
    if constexpr (T == 4 || T == 8)
        if (a.w->type == TB200_TYPE_Q4_0 && a.items && a.n_live && a.nmat <= 1 && a.world <= 1 && grouped_q4_0_smem(T, a.w->k) <= 110 * 1024)
            return launch_grouped_q4_0<T>(c, a);
    if (total < (long)c->sm_count * kWarps * 4) return launch_TR<T, 1>(c, a);
    // with peers the resident kernel's epilogue collects a CTA's 16 rows under ONE matrix: every matrix must then be whole CTAs
    bool cta_aligned = a.nrows % (kWarps * 2) == 0;
    for (int m = 1; m < a.nmat; ++m) cta_aligned = cta_aligned && a.w_more[m - 1]->rows % (kWarps * 2) == 0;
    if (a.w->type == TB200_TYPE_Q4_0 && !a.items && T <= 2 && resident_enabled() && total / 2 > 4 * 3L * c->sm_count * kWarps &&   // >= 4 rounds of the resident grid
        (size_t)T * (a.w->k / QK) * 40 * 3 <= 200 * 1024 && (a.world <= 1 || cta_aligned))
        return launch_resident_q4_0<T>(c, a);
    return launch_TR<T, 2>(c, a);
}

}  // namespace

int tb200_launch_gemv(tb200_ctx *c, const tb200_gemv_args &a) {
    if (a.nrows == 0 || a.tokens == 0) return TB200_OK;
    switch (a.tokens) {
    case 1: return launch_T<1>(c, a);
    case 2: return launch_T<2>(c, a);
    case 3: return launch_T<3>(c, a);
    case 4: return launch_T<4>(c, a);
    case 5: return launch_T<5>(c, a);
    case 6: return launch_T<6>(c, a);
    case 7: return launch_T<7>(c, a);
    case 8: return launch_T<8>(c, a);
    default:
        tb200_set_error("gemv: tokens=%ld exceeds TB200_GEMV_MAX_TOKENS", a.tokens);
        return TB200_ERR_INVALID;
    }
}

int tb200_launch_blockdot(tb200_ctx *c, const tb200_weights *w, const int8_t *aq, const int2 *ads, int32_t *out) {
    const long units = w->rows * (tb200_is_kquant(w->type) ? w->k / QK_K : w->k / QK);
    if (units == 0) return TB200_OK;
    blockdot_kernel<<<(unsigned)((units + 127) / 128), 128, 0, c->stream>>>(w->type, (const uint8_t *)w->qs, (const uint8_t *)w->sc, w->rows, w->k, aq, ads, out);
    tb200_count_launch();
    TB_CUDA(cudaGetLastError());
    return TB200_OK;
}
