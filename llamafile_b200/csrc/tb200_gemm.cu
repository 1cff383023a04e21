// Prefill path (SURVEY.md §8 rows a9-a12 at n > 8): quant tiles staged by TMA, dequantised to fp16 and fed to
// tcgen05.mma (cta_group::1, M=128, N<=256, K=16 per instruction) with fp32 accumulators in TMEM.
//
// Where the dequantised weights live: the first version wrote A' to shared memory (SS form of the MMA) and ran at
// ~980 clk per 64-wide K stage against 512 clk of tensor time: per stage the SM's shared memory had to move 16 KB of
// A' stores + 16 KB of A' reads + 32 KB of B' TMA fill + 32 KB of B' reads + the packed tile, ~104 KB = 203 B/clk
// against 128 B/clk. A' therefore goes to TENSOR MEMORY instead (tcgen05.st by the dequant warps, TS form of the
// MMA: A from TMEM, B from shared memory): a ring of kStages x 32 columns next to the accumulator. Shared memory then
// only carries B' and the packed tile (72 KB per stage, 140 B/clk). F16 weights need no dequantisation and keep the
// SS form (TMA writes the swizzled A tile).
//
// Arithmetic contract (DESIGN.md "prefill numerics"): the reference computes sum_b d_x[b]*d_y[b]*sum_j q_x*q_y with
// exact int32 block sums (llamafile/iqk_mul_mat.inc:1129-1201, tinyblas_cpu.h:934-993). Here both operands are
// turned into 16-bit floats once, A' = fp16(d_x*(q_x-8)) and B' = fp16(d_y*q_y*2^-e_t), and multiplied with fp32
// accumulation; each operand carries one rounding of 2^-12 relative (fp16 keeps 11 significand bits; bf16 would
// keep 8 and break the 1e-3 bound, which is why the 16-bit type is fp16 — same tensor-pipe rate). The activations
// are first quantized to Q8_0 / Q8_K bit-exactly like the reference, so quantization noise is identical and only
// the two operand roundings and the accumulation order differ. e_t is a per-token power of two that keeps every
// B' value inside fp16 range; it is undone exactly in the epilogue.
//
// K order: a dot product does not care about the order of k as long as A' and B' agree. A Q4_0 / Q4_K word holds
// nibbles of elements (j, j+16) / (l, l+32) interleaved, so unpacking it yields 8 halves in a fixed permuted order;
// the activation prep kernel writes B' in that same order, which saves all byte permutes in the dequant loop.
//
// Warp roles (384 threads, 1 CTA / SM): warp 0 TMA producer, warp 1 MMA issuer, warp 2 TMEM allocator,
// warps 4-11 dequantisers (and, after the main loop, TMEM->register->global epilogue). TMEM lane = weight row, so
// dequant warp w owns rows 32*(w&3).. of the stages i = (w-4)>>2 (mod 2).
#include <cuda.h>

#include "tb200_common.cuh"

namespace {

constexpr int BM = 128;        // weight rows per CTA tile = UMMA M
constexpr int BK = 64;         // K elements per stage = one 128-byte swizzle atom of fp16
constexpr int kDequantWarps = 16;
constexpr int kGroups = kDequantWarps / 4;   // dequant groups of four warps (one per 32 TMEM lanes); group g takes the stages i = g (mod kGroups)
// packed quant bytes of one weight row in one 64-wide K stage (0: F16 weights are not dequantised)
__host__ __device__ constexpr int packed_row_bytes(int type) {
    return (type == TB200_TYPE_Q8_0 || type == TB200_TYPE_G8) ? 64 : (type == TB200_TYPE_Q6_K ? 48 : (type == TB200_TYPE_F16 ? 0 : 32));
}
constexpr int kThreads = (4 + kDequantWarps) * 32;

// ------------------------------------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, uint64_t *bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
                 "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
// the same box written to the same shared-memory offset of every CTA in `mask`, its bytes counted on the barrier at the same offset there
__device__ __forceinline__ void tma_load_2d_mc(void *dst, const CUtensorMap *map, uint64_t *bar, int c0, int c1, uint16_t mask) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(dst)),
                 "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(mask) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap *map) { asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t *dst_smem, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// A from tensor memory (lane = row, 2 fp16 per 32-bit column, 8 columns per K=16 slice), B from shared memory
__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
        "}" ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
        "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]),
        "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// ... on the barrier at the same offset of every CTA in `mask` (B' tiles shared by a cluster: a stage is free when ALL its CTAs have read it)
__device__ __forceinline__ void umma_commit_mc(uint64_t *bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]),
          "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
          "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]),
          "=r"(v[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// ---- CTA-pair (cta_group::2) forms -------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa_rank(uint32_t smem_addr, uint32_t rank) {   // same offset in CTA `rank` of the cluster
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load into THIS CTA's shared memory whose bytes are counted on the barrier at `bar_cluster_addr` (the pair leader's)
__device__ __forceinline__ void tma_load_2d_pair(void *dst, const CUtensorMap *map, uint32_t bar_cluster_addr, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
                 "l"(map), "r"(bar_cluster_addr), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t *dst_smem, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t addr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
__device__ __forceinline__ void umma_f16_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
        "}" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_f16_ts_pair(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n"
        "}" ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// arrives (once all prior MMAs of this thread have retired) on the barrier at the same offset in both CTAs of the pair
__device__ __forceinline__ void umma_commit_pair(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)),
                 "h"((uint16_t)3) : "memory");
}

// shared-memory matrix descriptor, K-major, SWIZZLE_128B (cute::UMMA::SmemDescriptor bit layout): start>>4 [0,14),
// LBO>>4 [16,30) (ignored for swizzled K-major, canonical value 1), SBO>>4 [32,46) = 8 rows x 128 B = 1024 B,
// version=1 [46,48), layout_type=2 (SWIZZLE_128B) [61,64)
__device__ __forceinline__ uint64_t make_desc_sw128(uint32_t smem_addr) {
    return (uint64_t)((smem_addr & 0x3FFFF) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
// instruction descriptor (cute::UMMA::InstrDescriptor): c_format F32=1 [4,6), a/b_format F16=0 [7,10)/[10,13),
// a/b K-major = 0 [15]/[16], N>>3 [17,23), M>>4 [24,29)
__host__ __device__ constexpr uint32_t make_idesc_f16(int m, int n) { return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24); }

// ------------------------------------------------------------------------------------------------ dequantisers
__device__ __forceinline__ uint32_t h2_sub_mul(uint32_t x, __half2 sub, __half2 mul) {
    const __half2 r = __hmul2(__hsub2(*reinterpret_cast<const __half2 *>(&x), sub), mul);
    return *reinterpret_cast<const uint32_t *>(&r);
}
__device__ __forceinline__ uint32_t h2_fma(uint32_t x, __half2 mul, __half2 add) {
    const __half2 r = __hfma2(*reinterpret_cast<const __half2 *>(&x), mul, add);
    return *reinterpret_cast<const uint32_t *>(&r);
}

// Q4_0 word (bytes 4w..4w+3 of a block): halves come out as elements (4w, 4w+2, 4w+1, 4w+3, 4w+16, 4w+18, 4w+17, 4w+19).
// (nibble | 0x6400) is the fp16 1024+q; minus 1032 gives q-8 exactly; one rounding in the multiply by d.
__device__ __forceinline__ uint4 dequant_q4_word(uint32_t w, __half2 d2) {
    const __half2 k1032 = __floats2half2_rn(1032.f, 1032.f);
    uint4 o;
    o.x = h2_sub_mul((w & 0x000F000Fu) | 0x64006400u, k1032, d2);
    o.y = h2_sub_mul(((w >> 8) & 0x000F000Fu) | 0x64006400u, k1032, d2);
    o.z = h2_sub_mul(((w >> 4) & 0x000F000Fu) | 0x64006400u, k1032, d2);
    o.w = h2_sub_mul(((w >> 12) & 0x000F000Fu) | 0x64006400u, k1032, d2);
    return o;
}
// Q4_K word (bytes 4w..4w+3 of a 32-byte chunk): (l, l+2, l+1, l+3) of the low sub-block then of the high sub-block;
// value = (d*sc)*q - (dmin*m): fp16 fma of the exact small integer q with the two fp16-rounded sub-block constants.
__device__ __forceinline__ uint4 dequant_q4K_word(uint32_t w, __half2 a_lo, __half2 b_lo, __half2 a_hi, __half2 b_hi) {
    const __half2 k1024 = __floats2half2_rn(1024.f, 1024.f);
    auto q = [&](uint32_t x) { const __half2 r = __hsub2(*reinterpret_cast<const __half2 *>(&x), k1024); return *reinterpret_cast<const uint32_t *>(&r); };
    uint4 o;
    o.x = h2_fma(q((w & 0x000F000Fu) | 0x64006400u), a_lo, b_lo);
    o.y = h2_fma(q(((w >> 8) & 0x000F000Fu) | 0x64006400u), a_lo, b_lo);
    o.z = h2_fma(q(((w >> 4) & 0x000F000Fu) | 0x64006400u), a_hi, b_hi);
    o.w = h2_fma(q(((w >> 12) & 0x000F000Fu) | 0x64006400u), a_hi, b_hi);
    return o;
}
// Q8_0 word (4 int8): natural order. (byte ^ 0x80) | 0x6400 is 1024 + (q + 128); minus 1152 gives q exactly.
__device__ __forceinline__ uint2 dequant_q8_word(uint32_t w, __half2 d2) {
    const __half2 k1152 = __floats2half2_rn(1152.f, 1152.f);
    const uint32_t x = w ^ 0x80808080u;
    uint2 o;
    o.x = h2_sub_mul(__byte_perm(x, 0x64646464u, 0x4140), k1152, d2);
    o.y = h2_sub_mul(__byte_perm(x, 0x64646464u, 0x4342), k1152, d2);
    return o;
}

// G8 word (4 int8 of the generic form, tb200_common.cuh): value = a*q + b with the block's fp16 pair; q comes out exactly as for Q8_0.
__device__ __forceinline__ uint2 dequant_g8_word(uint32_t w, __half2 a2, __half2 b2) {
    const __half2 k1152 = __floats2half2_rn(1152.f, 1152.f);
    const uint32_t x = w ^ 0x80808080u;
    auto q = [&](uint32_t v) { const __half2 r = __hsub2(*reinterpret_cast<const __half2 *>(&v), k1152); return *reinterpret_cast<const uint32_t *>(&r); };
    uint2 o;
    o.x = h2_fma(q(__byte_perm(x, 0x64646464u, 0x4140)), a2, b2);
    o.y = h2_fma(q(__byte_perm(x, 0x64646464u, 0x4342)), a2, b2);
    return o;
}

__device__ __forceinline__ void q4k_scale_min(const uint4 hdr, int j, int &sc, int &m) {   // get_scale_min_k4, ggml-quants.c:1891-1899
    const uint32_t s0 = hdr.y, s1 = hdr.z, s2 = hdr.w;
    auto byte = [&](int i) -> uint32_t { const uint32_t wv = i < 4 ? s0 : (i < 8 ? s1 : s2); return (wv >> (8 * (i & 3))) & 0xff; };
    if (j < 4) { sc = byte(j) & 63; m = byte(j + 4) & 63; }
    else { sc = (byte(j + 4) & 0xF) | ((byte(j - 4) >> 6) << 4); m = (byte(j + 4) >> 4) | ((byte(j) >> 6) << 4); }
}

__host__ __device__ constexpr uint32_t tmem_cols_pow2(uint32_t need) {
    uint32_t c = 32;
    while (c < need) c <<= 1;
    return c;
}

struct GemmParams {
    const void *sc;        // scales plane of the weights
    long sc_row_stride;    // in elements of the plane (fp16 for Q4_0/Q8_0, uint4 for Q4_K / Q6_K)
    const void *d6;        // Q6_K: fp16 super-block scale plane, sc_row_stride elements per row
    long row0, nrows;      // weight rows [row0, row0 + nrows)
    long k, tokens;
    const float *colscale; // [tokens] 2^e_t
    float *out; long ldo;
    // grouped (mixmul) mode: blockIdx.y indexes a device-built tile table {weight row offset, first B' row, rows in tile, -};
    // rowmap[t] = output row of B' row t
    const int4 *tiles;
    const int *rowmap;
    long n_tiles;
    // fused all-gather (row-sharded prefill): `out` lies inside this rank's symmetric buffer `local_base`; every result is
    // stored at the same offset of every rank's copy (peer-mapped pointers, own entry included) over NVLink
    int world;
    const char *local_base;
    char *peer_base[TB200_MAX_PEERS];
};

// Epilogue store of one accumulator chunk: lane = weight row, so each of the 32 store instructions of a warp writes 128
// contiguous bytes of one token's output row. With world > 1 that segment goes to every rank's copy of the buffer: the
// NVLink transfers of a tile overlap the main loops of the other CTAs, and no collective follows — only the flag barrier.
__device__ __forceinline__ void store_chunk_out(const GemmParams &p, const uint32_t (&v)[32], long tile_tok, int ch, long tok_end, long out_row) {
    if (out_row >= p.nrows) return;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        const long t = tile_tok + ch * 32 + j;
        if (t < tok_end) {
            float *dst = p.out + (p.rowmap ? (long)p.rowmap[t] : t) * p.ldo + out_row;
            const float val = __uint_as_float(v[j]) * p.colscale[t];
            if (p.world <= 1) *dst = val;
            else {
                const long off = reinterpret_cast<const char *>(dst) - p.local_base;
                for (int pr = 0; pr < p.world; ++pr) *reinterpret_cast<float *>(p.peer_base[pr] + off) = val;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ dequant warps
// One dequant thread owns weight row `row` of the CTA tile (= its TMEM lane) for a whole 64-wide stage: two halves of
// 32 elements, each turned into 32 fp16 = 16 registers for one tcgen05.st.32x32b.x16. The 8 dequant warps form two
// groups of four (one warp per 32 TMEM lanes); group g takes the stages i = g (mod 2), so two stages are in flight and
// the ~400-clock chain wait -> ld.shared -> unpack -> tcgen05.st -> wait::st of one stage no longer bounds the rate
// (the first TS version had all 8 warps on the same stage and was dequant-latency-bound whenever a stage's MMAs took
// less than that, i.e. at BN <= 128). RowDequant carries the thread's scale stream: the scales of the group's next
// stage are fetched from global memory while the current one is being processed.
template <int kType>
struct RowDequant {
    const __half *scp;    // Q4_0 / Q8_0: fp16 d of this row; elements [2i], [2i+1] belong to the halves of stage i
    const uint4 *hdp;     // Q4_K: {d, dmin, scales[12]} of this row, one per 256 elements = 4 stages; Q6_K: int8 scales, word i = stage i's four
    const __half *d6p;    // Q6_K: fp16 d of this row, one per 256 elements
    // scale ring: entry j holds the scales of the group's stage (current + 2j). kLook group steps = 2*kLook stages of lookahead;
    // one step was enough while a stage's MMAs take >= 256 clocks (BN = 256) but left the L2/HBM latency of these scattered
    // 4..16-byte loads exposed at small token tiles, where a stage is ~100 clocks of tensor work (measured ~740 clocks per
    // stage at BN = 32 before this ring).
    static constexpr int kLook = 4;
    __half2 d_ring[kLook];
    uint4 h_ring[kLook];
    int row;

    __device__ __forceinline__ void fetch(int stage, __half2 &d, uint4 &h) const {
        if (kType == TB200_TYPE_G8) { const uint2 ab = __ldg(reinterpret_cast<const uint2 *>(scp) + stage); h.x = ab.x, h.y = ab.y; }   // {a, b} of the stage's two blocks
        else if (kType == TB200_TYPE_Q4_K) h = __ldg(hdp + (stage >> 2));
        else if (kType == TB200_TYPE_Q6_K) h.x = __ldg(reinterpret_cast<const uint32_t *>(hdp) + stage), d = __half2half2(d6p[stage >> 2]);
        else d = *reinterpret_cast<const __half2 *>(scp + 2 * stage);
    }
    __device__ __forceinline__ RowDequant(const GemmParams &p, long grow, int row_, int first_stage, int kt) : row(row_) {
        scp = reinterpret_cast<const __half *>(p.sc) + grow * p.sc_row_stride * (kType == TB200_TYPE_G8 ? 2 : 1);   // G8: 4 bytes per block
        hdp = reinterpret_cast<const uint4 *>(p.sc) + grow * p.sc_row_stride;
        d6p = reinterpret_cast<const __half *>(p.d6) + grow * p.sc_row_stride;
#pragma unroll
        for (int j = 0; j < kLook; ++j) {
            d_ring[j] = __half2half2(__float2half(0.f));
            h_ring[j] = make_uint4(0, 0, 0, 0);
            if (first_stage + kGroups * j < kt) fetch(first_stage + kGroups * j, d_ring[j], h_ring[j]);
        }
    }
    // call before waiting for stage i's packed tile: takes stage i's scales and puts the fetch for stage i + 2*kLook in flight
    __device__ __forceinline__ void advance(int i, int kt, __half2 &dpair, uint4 &hdr) {
        dpair = d_ring[0], hdr = h_ring[0];
#pragma unroll
        for (int j = 0; j + 1 < kLook; ++j) d_ring[j] = d_ring[j + 1], h_ring[j] = h_ring[j + 1];
        if (i + kGroups * kLook < kt) fetch(i + kGroups * kLook, d_ring[kLook - 1], h_ring[kLook - 1]);
    }
    // tile: this stage's packed quant tile in shared memory (TMA, 32B swizzle for 32-byte rows, 64B swizzle for 64-byte rows)
    __device__ __forceinline__ void regs(const uint8_t *tile, int i, int chalf, __half2 dpair, const uint4 &hdr, uint32_t (&r)[16]) const {
        const __half2 d2 = chalf ? __high2half2(dpair) : __low2half2(dpair);
        if (kType == TB200_TYPE_Q4_0) {
            // SWIZZLE_32B, 32-byte rows: 16-byte chunk c of row r sits at chunk c ^ ((r >> 2) & 1)
            const uint4 w = *reinterpret_cast<const uint4 *>(tile + row * 32 + ((chalf ^ ((row >> 2) & 1)) << 4));
            uint4 o;
            o = dequant_q4_word(w.x, d2), r[0] = o.x, r[1] = o.y, r[2] = o.z, r[3] = o.w;
            o = dequant_q4_word(w.y, d2), r[4] = o.x, r[5] = o.y, r[6] = o.z, r[7] = o.w;
            o = dequant_q4_word(w.z, d2), r[8] = o.x, r[9] = o.y, r[10] = o.z, r[11] = o.w;
            o = dequant_q4_word(w.w, d2), r[12] = o.x, r[13] = o.y, r[14] = o.z, r[15] = o.w;
        } else if (kType == TB200_TYPE_Q8_0) {
            // SWIZZLE_64B, 64-byte rows: chunk c of row r sits at chunk c ^ ((r >> 1) & 3)
            const uint8_t *rp = tile + row * 64;
            const int sw = (row >> 1) & 3;
            const uint4 w0 = *reinterpret_cast<const uint4 *>(rp + (((chalf * 2) ^ sw) << 4));
            const uint4 w1 = *reinterpret_cast<const uint4 *>(rp + (((chalf * 2 + 1) ^ sw) << 4));
            uint2 a;
            a = dequant_q8_word(w0.x, d2), r[0] = a.x, r[1] = a.y;
            a = dequant_q8_word(w0.y, d2), r[2] = a.x, r[3] = a.y;
            a = dequant_q8_word(w0.z, d2), r[4] = a.x, r[5] = a.y;
            a = dequant_q8_word(w0.w, d2), r[6] = a.x, r[7] = a.y;
            a = dequant_q8_word(w1.x, d2), r[8] = a.x, r[9] = a.y;
            a = dequant_q8_word(w1.y, d2), r[10] = a.x, r[11] = a.y;
            a = dequant_q8_word(w1.z, d2), r[12] = a.x, r[13] = a.y;
            a = dequant_q8_word(w1.w, d2), r[14] = a.x, r[15] = a.y;
        } else if (kType == TB200_TYPE_G8) {
            // same tile as Q8_0 (64-byte rows, SWIZZLE_64B); value = a*q + b with the fp16 pair of block 2i + chalf
            const uint8_t *rp = tile + row * 64;
            const int sw = (row >> 1) & 3;
            const uint4 w0 = *reinterpret_cast<const uint4 *>(rp + (((chalf * 2) ^ sw) << 4));
            const uint4 w1 = *reinterpret_cast<const uint4 *>(rp + (((chalf * 2 + 1) ^ sw) << 4));
            const uint32_t ab = chalf ? hdr.y : hdr.x;
            const __half2 abh = *reinterpret_cast<const __half2 *>(&ab);
            const __half2 a2 = __low2half2(abh), b2 = __high2half2(abh);
            uint2 a;
            a = dequant_g8_word(w0.x, a2, b2), r[0] = a.x, r[1] = a.y;
            a = dequant_g8_word(w0.y, a2, b2), r[2] = a.x, r[3] = a.y;
            a = dequant_g8_word(w0.z, a2, b2), r[4] = a.x, r[5] = a.y;
            a = dequant_g8_word(w0.w, a2, b2), r[6] = a.x, r[7] = a.y;
            a = dequant_g8_word(w1.x, a2, b2), r[8] = a.x, r[9] = a.y;
            a = dequant_g8_word(w1.y, a2, b2), r[10] = a.x, r[11] = a.y;
            a = dequant_g8_word(w1.z, a2, b2), r[12] = a.x, r[13] = a.y;
            a = dequant_g8_word(w1.w, a2, b2), r[14] = a.x, r[15] = a.y;
        } else if (kType == TB200_TYPE_Q6_K) {
            // stage i is unit u = i & 3 of super-block i >> 2: 48 bytes = pieces j = 0,1,2 (byte = q_j << 2 | two bits of q_3,
            // tb200_repack.cu); half `chalf` of the stage is j = 2*chalf, 2*chalf+1 (16 elements each, scale byte j of hdr.x):
            // value = fp16(d*sc_j) * (q - 32). A byte 4q under the fp16 exponent of 256 reads as 256 + q, so no shift is needed.
            const uint8_t *rp = tile + row * 48;     // 48-byte row stride: the 8 lanes of a quarter-warp hit 8 distinct bank groups
            const uint4 p0 = *reinterpret_cast<const uint4 *>(rp), p1 = *reinterpret_cast<const uint4 *>(rp + 16),
                        p2 = *reinterpret_cast<const uint4 *>(rp + 32);
            const uint32_t sw = chalf ? (hdr.x >> 16) : hdr.x;
            const int sA = (int)(int8_t)sw, sB = (int)(int8_t)(sw >> 8);
            const float d = __low2float(dpair);
            const __half2 aA = __float2half2_rn(d * (float)sA), aB = __float2half2_rn(d * (float)sB);
            const __half2 k288 = __floats2half2_rn(288.f, 288.f);
            const uint32_t a[4] = {p0.x, p0.y, p0.z, p0.w}, b[4] = {p1.x, p1.y, p1.z, p1.w}, h[4] = {p2.x, p2.y, p2.z, p2.w};
#pragma unroll
            for (int w = 0; w < 4; ++w) {
                const uint32_t xa = chalf ? (h[w] & 0xfcfcfcfcu) : (a[w] & 0xfcfcfcfcu);
                const uint32_t xb = chalf ? ((((a[w] << 6) & 0xc0c0c0c0u) | ((b[w] << 4) & 0x30303030u)) | ((h[w] << 2) & 0x0c0c0c0cu)) : (b[w] & 0xfcfcfcfcu);
                r[2 * w] = h2_sub_mul(__byte_perm(xa, 0x5c5c5c5cu, 0x4140), k288, aA);       // bytes -> fp16 256 + q, natural order
                r[2 * w + 1] = h2_sub_mul(__byte_perm(xa, 0x5c5c5c5cu, 0x4342), k288, aA);
                r[8 + 2 * w] = h2_sub_mul(__byte_perm(xb, 0x5c5c5c5cu, 0x4140), k288, aB);
                r[8 + 2 * w + 1] = h2_sub_mul(__byte_perm(xb, 0x5c5c5c5cu, 0x4342), k288, aB);
            }
        } else {   // Q4_K: stage i is 64-element chunk c = i & 3 of super-block i >> 2
            const __half2 dd = *reinterpret_cast<const __half2 *>(&hdr.x);
            const float d = __low2float(dd), dmin = __high2float(dd);
            int sc_lo, m_lo, sc_hi, m_hi;
            q4k_scale_min(hdr, 2 * (i & 3), sc_lo, m_lo);
            q4k_scale_min(hdr, 2 * (i & 3) + 1, sc_hi, m_hi);
            const __half2 a_lo = __float2half2_rn(d * sc_lo), b_lo = __float2half2_rn(-(dmin * m_lo));
            const __half2 a_hi = __float2half2_rn(d * sc_hi), b_hi = __float2half2_rn(-(dmin * m_hi));
            const uint4 w = *reinterpret_cast<const uint4 *>(tile + row * 32 + ((chalf ^ ((row >> 2) & 1)) << 4));
            uint4 o;
            o = dequant_q4K_word(w.x, a_lo, b_lo, a_hi, b_hi), r[0] = o.x, r[1] = o.y, r[2] = o.z, r[3] = o.w;
            o = dequant_q4K_word(w.y, a_lo, b_lo, a_hi, b_hi), r[4] = o.x, r[5] = o.y, r[6] = o.z, r[7] = o.w;
            o = dequant_q4K_word(w.z, a_lo, b_lo, a_hi, b_hi), r[8] = o.x, r[9] = o.y, r[10] = o.z, r[11] = o.w;
            o = dequant_q4K_word(w.w, a_lo, b_lo, a_hi, b_hi), r[12] = o.x, r[13] = o.y, r[14] = o.z, r[15] = o.w;
        }
    }
};

// ------------------------------------------------------------------------------------------------ the kernel
// kType: weight type. BN: token tile (UMMA N). kStages: pipeline depth.
// kC: CTAs of a cluster along the row tiles (1, 2 or 4) that work on the SAME token tile and share its B' stages: each loads
// 1/kC of the B' tile and multicasts it to all (B' is re-read once per 128 weight rows otherwise - at T = 4096, N = 14336 that is
// 3.7 GB of L2 reads per GEMM against 0.9 GB of weights, and the L2 -> SM fabric, ~6300 B/clk, is what bounds the kernel).
template <int kType, int BN, int kStages, int kC>
__global__ void __launch_bounds__(kThreads, 1) gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap map_w,
                                                                   const __grid_constant__ CUtensorMap map_b, const GemmParams p) {
    constexpr bool kTS = kType != TB200_TYPE_F16;   // A' in tensor memory (dequantised types) or in shared memory (F16, via TMA)
    constexpr int kPackedBytes = BM * packed_row_bytes(kType);
    constexpr int kABytes = kTS ? 0 : BM * BK * 2, kBBytes = BN * BK * 2;
    constexpr uint32_t kAccCols = BN < 32 ? 32 : BN;                 // accumulator columns [0, kAccCols)
    constexpr uint32_t kTmemCols = tmem_cols_pow2(kAccCols + (kTS ? kStages * 32 : 0));   // A' ring: 32 columns (64 fp16) per stage
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t *sA = smem;                                   // [kStages][BM x 128 B]  swizzled fp16 A (F16 weights only)
    uint8_t *sB = sA + kStages * kABytes;                 // [kStages][BN x 128 B]  swizzled fp16 B' (TMA)
    uint8_t *sP = sB + kStages * kBBytes;                 // [kStages][packed quant tile] (TMA, 32B/64B swizzle)
    uint64_t *bars = reinterpret_cast<uint64_t *>(sP + kStages * (kPackedBytes ? kPackedBytes : 16));
    uint64_t *full_p = bars, *full_a = bars + kStages, *full_b = bars + 2 * kStages, *empty = bars + 3 * kStages, *accum_full = bars + 4 * kStages;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(accum_full + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    long tile_row = (long)blockIdx.x * BM, tile_tok = (long)blockIdx.y * BN;
    long w_row0 = p.row0, tok_end = p.tokens;
    const int kt = (int)(p.k / BK);
    if (p.tiles) {   // CTA-uniform; the table is written by a predecessor kernel
        pdl_wait();
        const int4 tl = p.tiles[blockIdx.y];
        if (tl.z == 0) return;
        w_row0 = tl.x, tile_tok = tl.y, tok_end = tl.y + tl.z;
    }

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_w);
        tma_prefetch_desc(&map_b);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full_p[s], 1);
            mbar_init(&full_a[s], kType == TB200_TYPE_F16 ? 1 : 4);   // one elected arrive per warp of the stage's dequant group (F16: the TMA's expect_tx)
            mbar_init(&full_b[s], 1);
            mbar_init(&empty[s], kC);                    // one multicast commit per CTA of the cluster
        }
        mbar_init(accum_full, 1);
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_slot, kTmemCols);
    tc_fence_before();
    __syncthreads();
    if (kC > 1) cluster_sync_all();                      // every CTA's barriers exist before a peer's copy or commit can land on them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t crank = kC > 1 ? cluster_ctarank() : 0u;
    constexpr uint16_t kMask = (uint16_t)((1u << kC) - 1u);
    pdl_launch_dependents();
    pdl_wait();   // B' and colscale come from the activation prep kernel

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            for (int i = 0; i < kt; ++i) {
                const int s = i % kStages;
                if (i >= kStages) mbar_wait(&empty[s], ((i / kStages) - 1) & 1);
                if (kType == TB200_TYPE_F16) {
                    mbar_expect_tx(&full_a[s], kABytes);   // F16 weights: TMA writes the swizzled A tile directly
                    tma_load_2d(sA + s * kABytes, &map_w, &full_a[s], i * BK, (int)(w_row0 + tile_row));
                } else {
                    mbar_expect_tx(&full_p[s], kPackedBytes);
                    tma_load_2d(sP + s * kPackedBytes, &map_w, &full_p[s], i * packed_row_bytes(kType), (int)(w_row0 + tile_row));
                }
                mbar_expect_tx(&full_b[s], kBBytes);
                if (kC > 1) tma_load_2d_mc(sB + s * kBBytes + crank * (kBBytes / kC), &map_b, &full_b[s], i * BK, (int)tile_tok + (int)crank * (BN / kC), kMask);
                else tma_load_2d(sB + s * kBBytes, &map_b, &full_b[s], i * BK, (int)tile_tok);
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (one thread) =====
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_f16(BM, BN);
            for (int i = 0; i < kt; ++i) {
                const int s = i % kStages;
                const uint32_t ph = (i / kStages) & 1;
                mbar_wait(&full_a[s], ph);
                mbar_wait(&full_b[s], ph);
                tc_fence_after();
                const uint64_t db = make_desc_sw128(smem_u32(sB + s * kBBytes));
                if (kTS) {
                    const uint32_t ta = tmem_base + kAccCols + (uint32_t)s * 32;
#pragma unroll
                    for (int kk = 0; kk < BK / 16; ++kk)   // K=16 slice: 8 TMEM columns of A', 32 bytes inside B's swizzle atom
                        umma_f16_ts(tmem_base, ta + (uint32_t)(kk * 8), db + (uint64_t)(kk * 2), idesc, (i | kk) ? 1u : 0u);
                } else {
                    const uint64_t da = make_desc_sw128(smem_u32(sA + s * kABytes));
#pragma unroll
                    for (int kk = 0; kk < BK / 16; ++kk)
                        umma_f16(tmem_base, da + (uint64_t)(kk * 2), db + (uint64_t)(kk * 2), idesc, (i | kk) ? 1u : 0u);
                }
                if (kC > 1) umma_commit_mc(&empty[s], kMask);   // frees the stage in every CTA of the cluster when these MMAs have read it
                else umma_commit(&empty[s]);
            }
            umma_commit(accum_full);
        }
    } else if (warp >= 4) {
        // ===== dequantisers: packed tile (smem) -> fp16 -> tensor memory =====
        const int quad = warp & 3, grp = (warp - 4) >> 2;
        const int row = quad * 32 + lane;             // TMEM lane == weight row of the tile
        const long grow = w_row0 + min(tile_row + row, p.nrows - 1);   // clamp rows past the matrix end (their outputs are masked)
        if (kTS) {
            const uint32_t ta = tmem_base + ((uint32_t)(quad * 32) << 16) + kAccCols;
            RowDequant<kType> dq(p, grow, row, grp, kt);
            for (int i = grp; i < kt; i += kGroups) {   // this warp group's stages
                const int s = i % kStages;
                const uint32_t ph = (i / kStages) & 1;
                __half2 dpair = __half2half2(__float2half(0.f));
                uint4 hdr = make_uint4(0, 0, 0, 0);
                uint32_t r0[16], r1[16];
                dq.advance(i, kt, dpair, hdr);
                // A parity wait is only unambiguous one phase ahead. When the ring depth is not a multiple of the group count a
                // slot is served by a different group every turn, and this group could get here while the slot's previous
                // load (stage i - kStages, another group's) is still in flight — bulk copies may land out of order — and take
                // the phase before that one for its own. full_a[s] of the previous turn (that stage dequantised, hence landed)
                // is at most one phase behind here, because this group's last stage was only loaded after the in-order MMAs
                // had consumed everything up to i - kGroups - kStages. (Seen as a hang with 5 stages and 4 groups.)
                if (kStages % kGroups != 0 && i >= kStages) mbar_wait(&full_a[s], ph ^ 1);
                mbar_wait(&full_p[s], ph);
                dq.regs(sP + s * kPackedBytes, i, 0, dpair, hdr, r0);
                dq.regs(sP + s * kPackedBytes, i, 1, dpair, hdr, r1);
                // full_p[s] completing implies (through the producer's wait on empty[s]) that the MMAs which read this A' stage
                // one ring turn ago have retired
                tc_fence_after();
                tmem_st16(ta + (uint32_t)s * 32, r0);
                tmem_st16(ta + (uint32_t)s * 32 + 16, r1);
                tmem_wait_st();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&full_a[s]);
            }
        }
        // ===== epilogue: TMEM -> registers -> global. Warp w may touch TMEM lanes 32*(w%4).. ; warps 4-7 take the
        // first half of the columns, warps 8-11 the second half.
        mbar_wait(accum_full, 0);
        tc_fence_after();
        const long out_row = tile_row + quad * 32 + lane;
        constexpr int kChunks = BN / 32 > 0 ? BN / 32 : 1;
        for (int ch = grp; ch < kChunks; ch += kGroups) {
            uint32_t v[32];
            tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(ch * 32), v);
            store_chunk_out(p, v, tile_tok, ch, tok_end, out_row);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (kC > 1) cluster_sync_all();                      // no CTA leaves while a peer's commit may still arrive on its barriers
    if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
}

// ------------------------------------------------------------------------------------------------ the CTA-pair kernel
// Same pipeline on a pair of CTAs (cluster 2x1x1 along the row tiles, one UMMA of M=256): CTA r of the pair owns weight
// rows [128r, 128r+128) of a 256-row tile (its own packed tiles, its own dequant warps, its own TMEM lanes) and stages
// only HALF of the B' tile (tokens [r*BN/2, (r+1)*BN/2)); the tensor cores of both SMs read both halves. One SM's L2
// traffic per 64-wide stage drops from 4 KB + BN*128 B to 4 KB + BN*64 B, which is what the single-CTA kernel is bound
// by at BN=256 (36 KB per 512 tensor clocks per SM = 70 B/clk/SM against ~43 B/clk/SM of L2->SM fabric).
// Barrier ownership: full_a / full_b live in the leader (rank 0), where the one MMA thread waits — the peer's TMA
// counts its bytes there and the peer's dequant warps arrive there remotely; empty / accum_full exist in both CTAs
// and are arrived on by the leader's multicast tcgen05.commit; full_p is CTA-local.
template <int kType, int BN, int kStages>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
    gemm_tcgen05_pair_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_b, const GemmParams p) {
    constexpr bool kTS = kType != TB200_TYPE_F16;
    static_assert(!kTS || kStages % kGroups == 0, "every ring slot must belong to one dequant group (parity waits are one phase deep; full_a is remote for the peer)");
    constexpr int BNH = BN / 2;                      // B' rows staged by each CTA
    constexpr int kPackedBytes = BM * packed_row_bytes(kType);
    constexpr int kABytes = kTS ? 0 : BM * BK * 2, kBBytes = BNH * BK * 2;
    constexpr uint32_t kAccCols = BN;
    constexpr uint32_t kTmemCols = tmem_cols_pow2(kAccCols + (kTS ? kStages * 32 : 0));
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t *sA = smem;
    uint8_t *sB = sA + kStages * kABytes;
    uint8_t *sP = sB + kStages * kBBytes;
    uint64_t *bars = reinterpret_cast<uint64_t *>(sP + kStages * (kPackedBytes ? kPackedBytes : 16));
    uint64_t *full_p = bars, *full_a = bars + kStages, *full_b = bars + 2 * kStages, *empty = bars + 3 * kStages, *accum_full = bars + 4 * kStages;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(accum_full + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    long tile_row = (long)blockIdx.x * BM, tile_tok = (long)blockIdx.y * BN;
    long w_row0 = p.row0, tok_end = p.tokens;
    const int kt = (int)(p.k / BK);
    if (p.tiles) {   // uniform over the pair (same blockIdx.y); the table is written by a predecessor kernel
        pdl_wait();
        const int4 tl = p.tiles[blockIdx.y];
        if (tl.z == 0) return;
        w_row0 = tl.x, tile_tok = tl.y, tok_end = tl.y + tl.z;
    }

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_w);
        tma_prefetch_desc(&map_b);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full_p[s], 1);
            mbar_init(&full_a[s], kTS ? 8 : 1);   // the stage's dequant group (4 warps) in both CTAs (F16: the leader's expect_tx for both TMA halves)
            mbar_init(&full_b[s], 1);
            mbar_init(&empty[s], 1);
        }
        mbar_init(accum_full, 1);
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc_pair(tmem_slot, kTmemCols);
    tc_fence_before();
    cluster_sync_all();     // barriers of both CTAs initialised, both TMEM allocations done
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();
    pdl_wait();

    if (warp == 0) {
        // ===== TMA producer (one thread per CTA) =====
        if (lane == 0) {
            for (int i = 0; i < kt; ++i) {
                const int s = i % kStages;
                if (i >= kStages) mbar_wait(&empty[s], ((i / kStages) - 1) & 1);
                if (kType == TB200_TYPE_F16) {
                    if (leader) mbar_expect_tx(&full_a[s], 2 * kABytes);
                    tma_load_2d_pair(sA + s * kABytes, &map_w, mapa_rank(smem_u32(&full_a[s]), 0), i * BK, (int)(w_row0 + tile_row));
                } else {
                    mbar_expect_tx(&full_p[s], kPackedBytes);
                    tma_load_2d(sP + s * kPackedBytes, &map_w, &full_p[s], i * packed_row_bytes(kType), (int)(w_row0 + tile_row));
                }
                if (leader) mbar_expect_tx(&full_b[s], 2 * kBBytes);
                tma_load_2d_pair(sB + s * kBBytes, &map_b, mapa_rank(smem_u32(&full_b[s]), 0), i * BK, (int)(tile_tok + rank * BNH));
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: one thread of the leader drives the tensor cores of both SMs =====
        if (leader && lane == 0) {
            constexpr uint32_t idesc = make_idesc_f16(2 * BM, BN);
            for (int i = 0; i < kt; ++i) {
                const int s = i % kStages;
                const uint32_t ph = (i / kStages) & 1;
                mbar_wait(&full_a[s], ph);
                mbar_wait(&full_b[s], ph);
                tc_fence_after();
                const uint64_t db = make_desc_sw128(smem_u32(sB + s * kBBytes));
                if (kTS) {
                    const uint32_t ta = tmem_base + kAccCols + (uint32_t)s * 32;
#pragma unroll
                    for (int kk = 0; kk < BK / 16; ++kk)
                        umma_f16_ts_pair(tmem_base, ta + (uint32_t)(kk * 8), db + (uint64_t)(kk * 2), idesc, (i | kk) ? 1u : 0u);
                } else {
                    const uint64_t da = make_desc_sw128(smem_u32(sA + s * kABytes));
#pragma unroll
                    for (int kk = 0; kk < BK / 16; ++kk)
                        umma_f16_pair(tmem_base, da + (uint64_t)(kk * 2), db + (uint64_t)(kk * 2), idesc, (i | kk) ? 1u : 0u);
                }
                umma_commit_pair(&empty[s]);
            }
            umma_commit_pair(accum_full);
        }
    } else if (warp >= 4) {
        const int quad = warp & 3, grp = (warp - 4) >> 2;
        const int row = quad * 32 + lane;
        const long grow = w_row0 + min(tile_row + row, p.nrows - 1);
        if (kTS) {
            // ===== dequantisers: packed tile (smem) -> fp16 -> this CTA's tensor memory; arrive at the leader =====
            const uint32_t ta = tmem_base + ((uint32_t)(quad * 32) << 16) + kAccCols;
            RowDequant<kType> dq(p, grow, row, grp, kt);
            for (int i = grp; i < kt; i += kGroups) {   // this warp group's stages
                const int s = i % kStages;
                const uint32_t ph = (i / kStages) & 1;
                __half2 dpair = __half2half2(__float2half(0.f));
                uint4 hdr = make_uint4(0, 0, 0, 0);
                uint32_t r0[16], r1[16];
                dq.advance(i, kt, dpair, hdr);
                mbar_wait(&full_p[s], ph);
                dq.regs(sP + s * kPackedBytes, i, 0, dpair, hdr, r0);
                dq.regs(sP + s * kPackedBytes, i, 1, dpair, hdr, r1);
                // full_p[s] completing implies (through the producer's wait on empty[s]) that the MMAs which read this A' stage
                // one ring turn ago have retired
                tc_fence_after();
                tmem_st16(ta + (uint32_t)s * 32, r0);
                tmem_st16(ta + (uint32_t)s * 32 + 16, r1);
                tmem_wait_st();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    if (leader) mbar_arrive(&full_a[s]);
                    else mbar_arrive_cluster(mapa_rank(smem_u32(&full_a[s]), 0));
                }
            }
        }
        // ===== epilogue: this CTA's 128 rows x BN tokens =====
        mbar_wait(accum_full, 0);
        tc_fence_after();
        const long out_row = tile_row + quad * 32 + lane;
        constexpr int kChunks = BN / 32;
        for (int ch = grp; ch < kChunks; ch += kGroups) {
            uint32_t v[32];
            tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(ch * 32), v);
            store_chunk_out(p, v, tile_tok, ch, tok_end, out_row);
        }
    }
    tc_fence_before();
    cluster_sync_all();     // neither CTA leaves (or frees TMEM) while the other can still touch its barriers / memories
    if (warp == 2) tmem_dealloc_pair(tmem_base, kTmemCols);
}

// ------------------------------------------------------------------------------------------------ activation prep
// planar Q8 activations -> B' fp16 [tokens][k] in the weight type's K order, scaled by 2^-e_t; colscale[t] = 2^e_t.
// perm: 0 natural (Q8_0), 1 Q4_0 order per 32-block, 2 Q4_K order per 64-chunk, 3 Q6_K unit order per super-block.
__device__ __forceinline__ int perm_pos(int perm, int kk) {   // position of element kk inside its 64-element group
    if (perm == 0) return kk;
    if (perm == 1) {   // within a 32-block: element e -> word w = (e & 15) >> 2, slot: low nibble elems (4w,4w+2,4w+1,4w+3), high +16
        const int blk = kk >> 5, e = kk & 31, hi = e >> 4, x = e & 15, w = x >> 2, r = x & 3;
        const int slot = (r == 0 ? 0 : r == 2 ? 1 : r == 1 ? 2 : 3) + 4 * hi;
        return blk * 32 + w * 8 + slot;
    }
    const int hi = kk >> 5, x = kk & 31, w = x >> 2, r = x & 3;   // 64-chunk: low sub-block elems 0..31, high 32..63
    const int slot = (r == 0 ? 0 : r == 2 ? 1 : r == 1 ? 2 : 3) + 4 * hi;
    return w * 8 + slot;
}

// Grouped mode: B' row t is the activation row of pair order[t] (expert-sorted), i.e. row (pair / thinkers) * tasks + (pair % thinkers) % tasks.
__global__ void __launch_bounds__(256) prep_acts_kernel(const int8_t *__restrict__ aq, const int2 *__restrict__ ads, long a_stride, long k,
                                                        int perm, __half *__restrict__ bp, float *__restrict__ colscale,
                                                        const int *__restrict__ order, int thinkers, int tasks) {
    pdl_launch_dependents();
    pdl_wait();
    const long t = blockIdx.x;
    long src = t;
    if (order) { const int pr = order[t]; src = (long)(pr / thinkers) * tasks + (pr % thinkers) % tasks; }
    const int2 *ds = ads + src * (a_stride / QK);
    const int8_t *q = aq + src * a_stride;
    __shared__ float s_max[8];
    float bound = 0.f;
    for (long b = threadIdx.x; b < k / QK; b += blockDim.x) bound = fmaxf(bound, fabsf(__int_as_float(ds[b].x)) * 127.f);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) bound = fmaxf(bound, __shfl_xor_sync(0xffffffffu, bound, o));
    if ((threadIdx.x & 31) == 0) s_max[threadIdx.x >> 5] = bound;
    __syncthreads();
    bound = s_max[0];
#pragma unroll
    for (int i = 1; i < 8; ++i) bound = fmaxf(bound, s_max[i]);
    // e_t: bound * 2^-e in [2^14, 2^15) -> every |B'| < 65504; zero or non-finite rows keep e = 0
    int e = 0;
    if (bound > 0.f && bound < 3.0e38f) e = ilogbf(bound) - 14;
    const float inv = ldexpf(1.f, -e);
    if (threadIdx.x == 0) colscale[t] = ldexpf(1.f, e);
    // four consecutive elements per thread: one 32-bit load, one 8-byte store — every K order keeps a 4-run together
    // (perm 1/2: positions w*8 + 4*hi + {0,2,1,3}; perm 0/3: natural)
    const uint32_t *q4 = reinterpret_cast<const uint32_t *>(q);
    for (long i4 = threadIdx.x; i4 < k / 4; i4 += blockDim.x) {
        const long i = i4 * 4;
        const float d = __int_as_float(ds[i >> 5].x);
        const uint32_t qw = q4[i4];
        const float v0 = __fmul_rn(__fmul_rn(d, (float)(int8_t)(qw & 0xff)), inv), v1 = __fmul_rn(__fmul_rn(d, (float)(int8_t)((qw >> 8) & 0xff)), inv);
        const float v2 = __fmul_rn(__fmul_rn(d, (float)(int8_t)((qw >> 16) & 0xff)), inv), v3 = __fmul_rn(__fmul_rn(d, (float)(int8_t)(qw >> 24)), inv);
        long dst;
        __half2 lo, hi;
        if (perm == 3) {   // Q6_K: element 128*half + 32*j + 16*r + e of a super-block -> stage u = 2*half + r, position 16*j + e
            const int e256 = (int)(i & 255), hf = e256 >> 7, j = (e256 >> 5) & 3, r = (e256 >> 4) & 1, e = e256 & 15;
            dst = (i & ~255L) + (hf * 2 + r) * 64 + j * 16 + e;
            lo = __floats2half2_rn(v0, v1), hi = __floats2half2_rn(v2, v3);
        } else if (perm == 0) {
            dst = i;
            lo = __floats2half2_rn(v0, v1), hi = __floats2half2_rn(v2, v3);
        } else {
            dst = (i & ~63L) + perm_pos(perm, (int)(i & 63));   // position of the run's first element; the run is stored as (e0, e2, e1, e3)
            lo = __floats2half2_rn(v0, v2), hi = __floats2half2_rn(v1, v3);
        }
        uint2 o;
        o.x = *reinterpret_cast<const uint32_t *>(&lo), o.y = *reinterpret_cast<const uint32_t *>(&hi);
        *reinterpret_cast<uint2 *>(bp + t * k + dst) = o;
    }
}

__global__ void fill_ones_kernel(float *p, long n) {
    pdl_launch_dependents();
    pdl_wait();
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i < n) p[i] = 1.f;
}

// ------------------------------------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = [] {
        void *f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) f = nullptr;
        return (EncodeTiledFn)f;
    }();
    return fn;
}

bool make_map_2d(CUtensorMap *m, CUtensorMapDataType dt, size_t esize, void *base, uint64_t inner, uint64_t outer, uint64_t row_bytes,
                 uint32_t box_inner, uint32_t box_outer, CUtensorMapSwizzle sw) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) { tb200_set_error("gemm: cuTensorMapEncodeTiled unavailable"); return false; }
    const cuuint64_t dims[2] = {inner, outer};
    const cuuint64_t strides[1] = {row_bytes};
    const cuuint32_t box[2] = {box_inner, box_outer};
    const cuuint32_t estr[2] = {1, 1};
    (void)esize;
    const CUresult r = fn(m, dt, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { tb200_set_error("gemm: cuTensorMapEncodeTiled failed (%d)", (int)r); return false; }
    return true;
}

struct GemmScratch { void *p = nullptr; size_t cap = 0; };
GemmScratch g_scratch[16];

void *gemm_scratch(tb200_ctx *c, size_t bytes) {
    GemmScratch &s = g_scratch[c->device];
    if (bytes <= s.cap) return s.p;
    cudaStreamSynchronize(c->stream);
    if (s.p) cudaFree(s.p);
    s.p = nullptr, s.cap = 0;
    const size_t want = (bytes + bytes / 4 + (1 << 20) - 1) >> 20 << 20;
    if (cudaMalloc(&s.p, want) != cudaSuccess) { cudaGetLastError(); tb200_set_error("gemm: scratch cudaMalloc(%zu) failed", want); return nullptr; }
    s.cap = want;
    return s.p;
}

template <int kType, int BN, int kStages, int kC = 1>
int launch_cfg(tb200_ctx *c, const CUtensorMap &mw, const CUtensorMap &mb, const GemmParams &p) {
    constexpr int kPackedBytes = kType == TB200_TYPE_F16 ? 16 : BM * packed_row_bytes(kType);
    constexpr int kABytes = kType == TB200_TYPE_F16 ? BM * BK * 2 : 0;   // dequantised A' lives in tensor memory
    constexpr size_t smem = 1024 + (size_t)kStages * (kABytes + BN * BK * 2 + kPackedBytes) + (4 * kStages + 1) * 8 + 16;
    static_assert(smem <= 227 * 1024, "stage configuration exceeds shared memory");
    static_assert(kType == TB200_TYPE_F16 || (BN < 32 ? 32 : BN) + kStages * 32 <= 512, "accumulator + A' ring exceed tensor memory");
    static_assert(kC == 1 || (BN / kC) % 8 == 0, "a CTA's slice of the B' tile is whole 8-row swizzle groups");
    auto kern = gemm_tcgen05_kernel<kType, BN, kStages, kC>;
    static bool attr_set = false;
    if (!attr_set) {
        TB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set = true;
    }
    const unsigned row_tiles = (unsigned)((p.nrows + BM - 1) / BM);
    const dim3 grid((row_tiles + kC - 1) / kC * kC, p.tiles ? (unsigned)p.n_tiles : (unsigned)((p.tokens + BN - 1) / BN));   // a padding CTA computes masked rows
    if (kC == 1) {
        TB_CUDA(tb200_launch(kern, grid, dim3(kThreads), smem, c->stream, mw, mb, p));
        return TB200_OK;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid, cfg.blockDim = dim3(kThreads), cfg.dynamicSmemBytes = smem, cfg.stream = c->stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kC, attr[0].val.clusterDim.y = 1, attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr, cfg.numAttrs = tb200_pdl_enabled() ? 2 : 1;
    tb200_count_launch();
    TB_CUDA(cudaLaunchKernelEx(&cfg, kern, mw, mb, p));
    return TB200_OK;
}

// tensor map over the resident quants plane (whole matrix), cached on the weights object. Packed tiles use the 32-byte
// (Q4_0/Q4_K: 32 bytes per row and stage) or 64-byte (Q8_0) swizzle so that the dequant warps' row-strided 16-byte
// reads are bank-conflict free.
CUtensorMap *weight_map(const tb200_weights *w) {
    if (!w->tmap_cache) {
        CUtensorMap *m = new CUtensorMap();
        bool ok;
        if (w->type == TB200_TYPE_F16) ok = make_map_2d(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, w->qs, (uint64_t)w->k, (uint64_t)w->rows, (uint64_t)w->k * 2, BK, BM, CU_TENSOR_MAP_SWIZZLE_128B);
        else if (w->type == TB200_TYPE_Q8_0 || tb200_is_g8(w->type)) ok = make_map_2d(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, w->qs, (uint64_t)w->k, (uint64_t)w->rows, (uint64_t)w->k, BK, BM, CU_TENSOR_MAP_SWIZZLE_64B);
        else if (w->type == TB200_TYPE_Q6_K) ok = make_map_2d(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, w->qs, (uint64_t)w->k / QK_K * 192, (uint64_t)w->rows, (uint64_t)w->k / QK_K * 192, 48, BM, CU_TENSOR_MAP_SWIZZLE_NONE);
        else ok = make_map_2d(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, w->qs, (uint64_t)w->k / 2, (uint64_t)w->rows, (uint64_t)w->k / 2, BK / 2, BM, CU_TENSOR_MAP_SWIZZLE_32B);
        if (!ok) { delete m; return nullptr; }
        const_cast<tb200_weights *>(w)->tmap_cache = m;
    }
    return reinterpret_cast<CUtensorMap *>(w->tmap_cache);
}

template <int kType, int BN, int kStages>
int launch_cfg_pair(tb200_ctx *c, const CUtensorMap &mw, const CUtensorMap &mb, const GemmParams &p) {
    constexpr int kPackedBytes = kType == TB200_TYPE_F16 ? 16 : BM * packed_row_bytes(kType);
    constexpr int kABytes = kType == TB200_TYPE_F16 ? BM * BK * 2 : 0;
    constexpr size_t smem = 1024 + (size_t)kStages * (kABytes + (BN / 2) * BK * 2 + kPackedBytes) + (4 * kStages + 1) * 8 + 16;
    static_assert(smem <= 227 * 1024, "stage configuration exceeds shared memory");
    static_assert(kType == TB200_TYPE_F16 || BN + kStages * 32 <= 512, "accumulator + A' ring exceed tensor memory");
    auto kern = gemm_tcgen05_pair_kernel<kType, BN, kStages>;
    static bool attr_set = false;
    if (!attr_set) {
        TB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set = true;
    }
    const unsigned row_tiles = (unsigned)((p.nrows + BM - 1) / BM);
    const dim3 grid((row_tiles + 1) & ~1u, p.tiles ? (unsigned)p.n_tiles : (unsigned)((p.tokens + BN - 1) / BN));   // pairs along x
    TB_CUDA(tb200_launch(kern, grid, dim3(kThreads), smem, c->stream, mw, mb, p));
    return TB200_OK;
}

// TB200_GEMM_PAIR: 0 = never, 1 (default) = CTA pairs where measured faster, 2 = CTA pairs for every weight type.
// Measured on B200 (tools/gemm_probe.py, N=14336 K=4096 T=4096, sustained): F16 weights (SS form) 792 -> 841 TFLOPS with
// pairs; the dequantised types (TS form) run at 864 TFLOPS single-CTA but only 518 TFLOPS as pairs (not power-limited
// there: 810 W, full clock — a latency stall in the pair's A'-in-TMEM hand-off that is not understood yet), so they
// stay on the single-CTA kernel.
int pair_mode() {
    static const int mode = [] { const char *v = getenv("TB200_GEMM_PAIR"); return (v && *v >= '0' && *v <= '2') ? *v - '0' : 1; }();
    return mode;
}

// deepest ring that fits: shared memory (227 KB minus alignment and barriers) and, for the TS form, tensor memory
constexpr int deep_stages(int type, int bn) {
    const int stage = (type == TB200_TYPE_F16 ? BM * BK * 2 : BM * packed_row_bytes(type)) + bn * BK * 2;
    int s = (227 * 1024 - 2048) / (stage + 32);
    if (type != TB200_TYPE_F16 && s > (512 - bn) / 32) s = (512 - bn) / 32;
    return s > 16 ? 16 : s;
}

template <int kType>
int launch_type(tb200_ctx *c, const tb200_weights *w, CUtensorMap &mw, long tokens, void *bp, GemmParams &p, int force_bn = 0) {
    // token tile: as wide as the batch allows (fewer re-dequantisations of the same weight tile), but keep >= ~1 wave of CTAs
    const long row_tiles = (p.nrows + BM - 1) / BM;
    int bn = 256;
    while (bn > 32 && (tokens <= bn / 2 || row_tiles * ((tokens + bn - 1) / bn) < c->sm_count)) bn >>= 1;
    if (tokens > 128 && bn < 128) bn = 128;
    if (force_bn) bn = force_bn;
    const bool pair = bn >= 128 && (pair_mode() == 2 || (pair_mode() == 1 && kType == TB200_TYPE_F16));   // CTA pairs: each CTA stages half of the B' tile
    // clusters of row-tile CTAs that multicast their slices of the B' tile to each other (TB200_GEMM_CLUSTER=2|4). OFF by default:
    // parity-green, but measured equal or slower on every BASELINE shape (Q8_0 14336x4096, T=4096: 513 / 523 / 556 us for cluster
    // sizes 1 / 2 / 4) - multicast to <= 4 CTAs does not reduce what the L2 delivers (B300_MICROARCH: "At csz<=4, MC = UC").
    static const int cluster_env = [] { const char *v = getenv("TB200_GEMM_CLUSTER"); return v && *v ? atoi(v) : 1; }();
    const int kc = (!pair && !p.tiles && bn >= 128 && row_tiles >= 2 && (cluster_env == 2 || cluster_env == 4)) ? cluster_env : 1;
    CUtensorMap mb;
    if (!make_map_2d(&mb, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, bp, (uint64_t)w->k, (uint64_t)tokens, (uint64_t)w->k * 2, BK, (uint32_t)(pair ? bn / 2 : bn / kc), CU_TENSOR_MAP_SWIZZLE_128B))
        return TB200_ERR_CUDA;
    constexpr bool f16 = kType == TB200_TYPE_F16;   // SS form: A and B stages in shared memory; TS form: B + packed tile only
    if (pair) {
        if (bn == 256) return launch_cfg_pair<kType, 256, f16 ? 6 : 8>(c, mw, mb, p);
        return launch_cfg_pair<kType, 128, 8>(c, mw, mb, p);
    }
    constexpr int kS256 = f16 ? 4 : ((kType == TB200_TYPE_Q8_0 || kType == TB200_TYPE_Q6_K || kType == TB200_TYPE_G8) ? 5 : 6);
    if (kc == 2) return bn == 256 ? launch_cfg<kType, 256, kS256, 2>(c, mw, mb, p) : launch_cfg<kType, 128, deep_stages(kType, 128), 2>(c, mw, mb, p);
    if (kc == 4) return bn == 256 ? launch_cfg<kType, 256, kS256, 4>(c, mw, mb, p) : launch_cfg<kType, 128, deep_stages(kType, 128), 4>(c, mw, mb, p);
    switch (bn) {
    case 256: return launch_cfg<kType, 256, kS256>(c, mw, mb, p);
    // small token tiles are weight-streaming-bound: the ring is as deep as tensor memory (accumulator + 32 columns of A' per
    // stage <= 512) and shared memory allow, since bytes in flight per SM are what sets the HBM rate there
    case 128: return launch_cfg<kType, 128, deep_stages(kType, 128)>(c, mw, mb, p);
    case 64: return launch_cfg<kType, 64, deep_stages(kType, 64)>(c, mw, mb, p);
    default: return launch_cfg<kType, 32, deep_stages(kType, 32)>(c, mw, mb, p);
    }
}


// ---- grouped (mixmul) support ------------------------------------------------------------------------------------
// One CTA; warp e scans all (token, thinker) pairs in order (same stable order as the reference's build_row_pointers,
// tinyblas_cpu_mixmul.inc:297-320): order[] = pairs sorted by expert, tiles[] = {expert row offset, first sorted row, count}.
__global__ void __launch_bounds__(1024) build_groups_kernel(const int32_t *__restrict__ plan, long pairs, int experts, long rows, int bn,
                                                            int *__restrict__ order, int4 *__restrict__ tiles, long max_tiles) {
    extern __shared__ int s_cnt[];   // [experts + 1]
    pdl_launch_dependents();
    pdl_wait();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    for (int e = warp; e < experts; e += nwarps) {
        int count = 0;
        for (long p0 = 0; p0 < pairs; p0 += 32) {
            const long pr = p0 + lane;
            count += __popc(__ballot_sync(0xffffffffu, pr < pairs && plan[pr] == e));
        }
        if (lane == 0) s_cnt[e + 1] = count;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        s_cnt[0] = 0;
        long nt = 0;
        for (int e = 0; e < experts; ++e) {
            const int cnt = s_cnt[e + 1], off = s_cnt[e];
            for (int t0 = 0; t0 < cnt && nt < max_tiles; t0 += bn) tiles[nt++] = make_int4((int)(e * rows), off + t0, min(bn, cnt - t0), e);
            s_cnt[e + 1] = off + cnt;
        }
        for (; nt < max_tiles; ++nt) tiles[nt] = make_int4(0, 0, 0, 0);
    }
    __syncthreads();
    // sorted slots past the routed pairs (plan entries naming no expert) belong to no tile; point them at pair 0 so the
    // activation prep reads a valid row
    for (long i = s_cnt[experts] + threadIdx.x; i < pairs; i += blockDim.x) order[i] = 0;
    for (int e = warp; e < experts; e += nwarps) {
        int count = 0;
        for (long p0 = 0; p0 < pairs; p0 += 32) {
            const long pr = p0 + lane;
            const bool hit = pr < pairs && plan[pr] == e;
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            if (hit) order[s_cnt[e] + count + __popc(m & ((1u << lane) - 1))] = (int)pr;
            count += __popc(m);
        }
    }
}

__global__ void gather_f16_kernel(const __half *__restrict__ a, long k, const int *__restrict__ order, int thinkers, int tasks,
                                  __half *__restrict__ bp, float *__restrict__ colscale) {
    pdl_launch_dependents();
    pdl_wait();
    const long t = blockIdx.x;
    const int pr = order[t];
    const long src = (long)(pr / thinkers) * tasks + (pr % thinkers) % tasks;
    for (long i = threadIdx.x; i < k; i += blockDim.x) bp[t * k + i] = a[src * k + i];
    if (threadIdx.x == 0) colscale[t] = 1.f;
}

}  // namespace

// mixmul prefill: all experts in one grouped tcgen05 launch. Activations: planar (aq, ads) [tokens*tasks][k] or fp16 (af).
// result[pair][rows]. Pairs whose plan entry names no expert are in no tile and stay untouched.
int tb200_launch_gemm_grouped(tb200_ctx *c, const tb200_weights *w, int experts, long rows, const int8_t *aq, const int2 *ads, const void *af,
                              const int32_t *plan, long pairs, int thinkers, int tasks, float *result) {
    if (w->k % 256 || ((uintptr_t)w->qs & 15)) { tb200_set_error("gemm: K=%ld must be a multiple of 256", w->k); return TB200_ERR_UNSUPPORTED; }
    const long avg = pairs / experts;
    const int bn = avg >= 192 ? 256 : (avg >= 96 ? 128 : (avg >= 40 ? 64 : 32));
    const long max_tiles = (pairs + bn - 1) / bn + experts;
    const size_t bp_bytes = ((size_t)(pairs + 256) * w->k * 2 + 255) & ~(size_t)255;   // + one box of slack rows for the last tile's TMA
    const size_t cs_bytes = ((size_t)pairs * 4 + 255) & ~(size_t)255, ord_bytes = cs_bytes, tl_bytes = ((size_t)max_tiles * 16 + 255) & ~(size_t)255;
    uint8_t *scr = (uint8_t *)gemm_scratch(c, bp_bytes + cs_bytes + ord_bytes + tl_bytes);
    if (!scr) return TB200_ERR_CUDA;
    __half *bp = (__half *)scr;
    float *colscale = (float *)(scr + bp_bytes);
    int *order = (int *)(scr + bp_bytes + cs_bytes);
    int4 *tiles = (int4 *)(scr + bp_bytes + cs_bytes + ord_bytes);
    const int threads = experts >= 32 ? 1024 : experts * 32;
    TB_CUDA(tb200_launch(build_groups_kernel, dim3(1), dim3(threads), (experts + 1) * sizeof(int), c->stream, plan, pairs, experts, rows, bn, order, tiles, max_tiles));
    // rows of B' that belong to no expert (invalid plan entries) are never read by a tile; order[] past the valid count is
    // not written, so size the prep grid by a device-independent bound: every sorted slot < pairs is either valid or unused.
    // Unused slots would read garbage order[] -> pre-fill with pair 0.
    if (w->type == TB200_TYPE_F16) {
        TB_CUDA(tb200_launch(gather_f16_kernel, dim3((unsigned)pairs), dim3(256), 0, c->stream, (const __half *)af, w->k, (const int *)order, thinkers, tasks, bp, colscale));
    } else {
        const int perm = w->type == TB200_TYPE_Q8_0 ? 0 : (w->type == TB200_TYPE_Q4_0 ? 1 : (w->type == TB200_TYPE_Q6_K ? 3 : 2));
        TB_CUDA(tb200_launch(prep_acts_kernel, dim3((unsigned)pairs), dim3(256), 0, c->stream, aq, ads, w->k, w->k, perm, bp, colscale, (const int *)order, thinkers, tasks));
    }
    CUtensorMap *mwp = weight_map(w);
    if (!mwp) return TB200_ERR_CUDA;
    CUtensorMap &mw = *mwp;
    GemmParams p;
    p.sc = w->sc;
    p.sc_row_stride = tb200_is_kquant(w->type) ? w->k / QK_K : w->k / QK;
    p.d6 = w->d;
    p.row0 = 0, p.nrows = rows, p.k = w->k, p.tokens = pairs, p.colscale = colscale, p.out = result, p.ldo = rows;
    p.tiles = tiles, p.rowmap = order, p.n_tiles = max_tiles;
    p.world = 1, p.local_base = nullptr;
    for (int r = 0; r < TB200_MAX_PEERS; ++r) p.peer_base[r] = nullptr;
    switch (w->type) {
    case TB200_TYPE_Q4_0: return launch_type<TB200_TYPE_Q4_0>(c, w, mw, pairs, bp, p, bn);
    case TB200_TYPE_Q8_0: return launch_type<TB200_TYPE_Q8_0>(c, w, mw, pairs, bp, p, bn);
    case TB200_TYPE_F16: return launch_type<TB200_TYPE_F16>(c, w, mw, pairs, bp, p, bn);
    default: return TB200_ERR_UNSUPPORTED;
    }
}

namespace {
}  // namespace

bool tb200_gemm_available() {
    static const int on = [] { const char *v = getenv("TB200_GEMM"); return (v && *v == '0') ? 0 : 1; }();
    return on != 0 && encode_fn() != nullptr;
}

void tb200_gemm_release(tb200_weights *w) {
    if (w && w->tmap_cache) { delete reinterpret_cast<CUtensorMap *>(w->tmap_cache); w->tmap_cache = nullptr; }
}

// out[t*ldo + i] for rows [row0, row0+nrows). Quantized types take planar activations (aq, ads); F16 takes af (fp16 [tokens][a_stride]).
int tb200_launch_gemm(tb200_ctx *c, const tb200_weights *w, long row0, long nrows, const int8_t *aq, const int2 *ads, const void *af,
                      long a_stride, long tokens, float *out, long ldo, const tb200_peers *peers) {
    if (w->k % 256 || ((uintptr_t)w->qs & 15)) { tb200_set_error("gemm: K=%ld must be a multiple of 256", w->k); return TB200_ERR_UNSUPPORTED; }
    const size_t bp_bytes = ((size_t)tokens * w->k * 2 + 255) & ~(size_t)255;
    uint8_t *scr = (uint8_t *)gemm_scratch(c, bp_bytes + (size_t)tokens * 4);
    if (!scr) return TB200_ERR_CUDA;
    __half *bp = (__half *)scr;
    float *colscale = (float *)(scr + bp_bytes);
    const void *b_src = bp;
    if (w->type == TB200_TYPE_F16) {
        // ggml already rounded the activations to fp16 (ggml_fp32_to_fp16_row): use them as they are
        if (a_stride != w->k) { tb200_set_error("gemm: F16 activations must be contiguous"); return TB200_ERR_UNSUPPORTED; }
        b_src = af;
        TB_CUDA(tb200_launch(fill_ones_kernel, dim3((unsigned)((tokens + 255) / 256)), dim3(256), 0, c->stream, colscale, tokens));
    } else {
        const int perm = (w->type == TB200_TYPE_Q8_0 || tb200_is_g8(w->type)) ? 0 : (w->type == TB200_TYPE_Q4_0 ? 1 : (w->type == TB200_TYPE_Q6_K ? 3 : 2));
        TB_CUDA(tb200_launch(prep_acts_kernel, dim3((unsigned)tokens), dim3(256), 0, c->stream, aq, ads, a_stride, w->k, perm, bp, colscale, (const int *)nullptr, 1, 1));
    }
    // weight tensor map (whole matrix), cached on the weights object
    CUtensorMap *mwp = weight_map(w);
    if (!mwp) return TB200_ERR_CUDA;
    CUtensorMap &mw = *mwp;
    GemmParams p;
    p.sc = w->sc;
    p.sc_row_stride = (tb200_is_kquant(w->type) && !tb200_is_g8(w->type)) ? w->k / QK_K : w->k / QK;
    p.d6 = w->d;
    p.row0 = row0, p.nrows = nrows, p.k = w->k, p.tokens = tokens, p.colscale = colscale, p.out = out, p.ldo = ldo;
    p.tiles = nullptr, p.rowmap = nullptr, p.n_tiles = 0;
    p.world = 1, p.local_base = nullptr;
    for (int r = 0; r < TB200_MAX_PEERS; ++r) p.peer_base[r] = nullptr;
    if (peers && peers->world > 1) {
        p.world = peers->world, p.local_base = (const char *)peers->out_base[peers->rank];
        for (int r = 0; r < peers->world; ++r) p.peer_base[r] = (char *)peers->out_base[r];
    }
    switch (w->type) {
    case TB200_TYPE_Q4_0: return launch_type<TB200_TYPE_Q4_0>(c, w, mw, tokens, (void *)b_src, p);
    case TB200_TYPE_Q8_0: return launch_type<TB200_TYPE_Q8_0>(c, w, mw, tokens, (void *)b_src, p);
    case TB200_TYPE_Q4_K: return launch_type<TB200_TYPE_Q4_K>(c, w, mw, tokens, (void *)b_src, p);
    case TB200_TYPE_Q6_K: return launch_type<TB200_TYPE_Q6_K>(c, w, mw, tokens, (void *)b_src, p);
    case TB200_TYPE_F16: return launch_type<TB200_TYPE_F16>(c, w, mw, tokens, (void *)b_src, p);
    case TB200_TYPE_Q4_1: case TB200_TYPE_Q5_0: case TB200_TYPE_Q5_K: case TB200_TYPE_IQ4_XS: return launch_type<TB200_TYPE_G8>(c, w, mw, tokens, (void *)b_src, p);
    default: return TB200_ERR_UNSUPPORTED;
    }
}
