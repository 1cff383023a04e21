// Decode plan: ONE persistent kernel for a whole sequence of decode (one-token) matmul groups.
//
// Why. A decode step of an 8B model is 129 launch groups (wq/wk/wv, wo, gate/up, down per layer + the vocabulary matrix)
// of 9..66 MB each, i.e. 1.5..10 us of HBM time per group. As one-shot launches (tb200_gemv.cu) every group pays a launch
// ramp, a drain and a wave tail (~1.5 us, more than the data time of the small groups), and the activation quantizer is
// a second launch per group: the step ran at 0.57 of the HBM roofline although each kernel moves every byte exactly once.
// Here the grid is one CTA per SM for the whole step, and the weight stream never stops at a group boundary:
//
//   * every warp owns a private ring of shared-memory stages that it fills itself with 1-D TMA bulk copies
//     (cp.async.bulk.shared.global + mbarrier complete_tx): a stage holds one 128-block K segment of a ROW PAIR (4 KB of
//     nibbles + 512 B of fp16 scales). The warp's work list - (group, row pair, segment) - is a pure function of the plan,
//     and resident weights are immutable, so the copy for item i+S is issued as soon as item i has been consumed, ACROSS
//     group boundaries: while a warp waits for the previous group to finish, its ring already holds the next group's rows,
//     and ~150 KB per SM (22 MB per GPU) stay in flight without costing a register;
//   * between groups there is one grid-wide barrier (arrival counter in L2) because group g's input is, in a model, a
//     function of group g-1's output; the activations of the new group are then produced INSIDE the kernel by every CTA
//     (16 KB of fp32 from L2, bit-exact Q8_0 quantization into shared memory), so the separate quantizer launch is gone;
//   * the neighbours of the matmul in the reference graph are fused into that prologue / the epilogue:
//     RMSNorm * weight -> Q8_0 (llm_build_norm + from_float, llama.cpp/src/llama.cpp:8099-8225), silu(gate) * up -> Q8_0
//     (llm_build_ffn, LLM_FFN_SILU + LLM_FFN_PAR) and the residual add on the way out.
//
// Arithmetic per block is the decode kernels' (ggml_vec_dot_q4_0_q8_0, ggml-quants.c:4192-4206): exact int32 block sums
// by dp4a, -8 folded in as -8*sum(y), fp32 scale FMA per block, warp-shuffle reduction per row.
#include <stdlib.h>
#include <string.h>

#include "tb200_common.cuh"

namespace {

#ifndef TB200_PLAN_WARPS
#define TB200_PLAN_WARPS 15
#endif
#ifndef TB200_PLAN_STAGES
#define TB200_PLAN_STAGES 2
#endif
#ifndef TB200_PLAN_ROWS
#define TB200_PLAN_ROWS 2
#endif
constexpr int kPW = TB200_PLAN_WARPS;                     // consumer warps per CTA, each with its own ring
constexpr int kTW = kPW + 1;                              // + the sync warp, which runs the grid-barrier protocol (and helps staging the activations)
constexpr int kPS = TB200_PLAN_STAGES;                    // ring stages per warp
constexpr int kR = TB200_PLAN_ROWS;                       // weight rows per work unit (they share every activation read)
constexpr int kSegBlocks = 128;                           // 32-element blocks of K per stage and row (4096 elements)
constexpr int kRowQ = kSegBlocks * 16, kRowS = kSegBlocks * 2;
constexpr int kStageBytes = kR * (kRowQ + kRowS);         // kR rows x (2 KB of nibbles + 256 B of fp16 scales): 8 x 2 x 9 KB = 147 KB in flight per SM
constexpr int kPlanMats = TB200_GEMV_MAX_MATS;
constexpr bool kActCache = kTW * 32 <= 384;               // activations of K <= 4096 groups in registers (needs > 128 registers per thread)

struct alignas(16) PlanGroup {
    const float *x, *x2, *nw;          // prologue inputs (see tb200_plan_group)
    float eps;
    int prologue;
    int nb, nseg;                      // 32-blocks per row; ring stages per unit
    int nunits, rot;                   // units (kR consecutive rows of one matrix) of all matrices; rotation of the warp -> unit assignment
    int nmat;
    int nodep;                         // the caller declared the group independent of earlier groups of the launch (TB200_PLAN_INDEPENDENT): no grid barrier before it
    int pad_;
    int shared_mask, sh_before;        // multi-GPU plans: bit m = matrix m's rows are stored into every rank's symmetric buffer; shared groups before this one
    int unit_begin[kPlanMats + 1];
    int rows[kPlanMats];
    const uint4 *qs[kPlanMats];
    const __half *sc[kPlanMats];
    float *out[kPlanMats];
    const float *res[kPlanMats];
};

static_assert(sizeof(PlanGroup) % 16 == 0, "descriptors are copied to shared memory in 16-byte words");
constexpr size_t kPlanSmemMax = 227 * 1024 - 512;         // dynamic shared memory the kernel may be given (227 KB per CTA minus its static part, < 400 B)

struct PlanParams {
    const PlanGroup *groups;
    int first, end;                    // groups [first, end) run in this launch
    unsigned *cnt;                     // [n_groups] arrival counters, all zero between launches
    int *err;                          // set when a wait gave up (the results of that launch are then undefined)
    int nb_max;
    int flags;                         // experiments (TB200_PLAN_FLAGS): 2 = stage activations only for the first group, 4 = no grid barrier (both wrong results),
                                       // 8 = plain result stores + release fence instead of confirmed exchanges (1 with 8: no fence, unsafe)
    long long *trace;                  // optional (tb200_plan_trace): [group][CTA][8] SM-clock stamps of the group's phases
    // multi-GPU plans (tb200_plan_create_peers): one process per GPU runs the same group sequence on its row shards
    int world, rank, n_shared;         // n_shared: shared groups of the whole plan
    const char *local_base;            // this rank's symmetric buffer; a shared result at local_base + off also goes to peer_base[p] + off
    char *peer_base[TB200_MAX_PEERS];
    unsigned *peer_flag[TB200_MAX_PEERS];   // rank p's flag words: [kPeerFlag0 + r] = the last shared-group sequence number rank r has completed
    unsigned *epoch, *done;            // the peer group's count of completed shared groups (tb200_peers::plan_seq); CTAs that left the launch
};
constexpr int kPeerFlag0 = 8;           // words 0..7 of a rank's flag buffer belong to tb200_peers_barrier

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
// mbarrier operations and the 1-D bulk copy global -> shared (completion counted in bytes on an mbarrier), on precomputed 32-bit
// shared-memory addresses: the ring's refill path runs once per 4.6 KB item. The weights are read exactly once, so the copies mark
// their lines evict-first in L2 and leave the activations / outputs of the step resident there.
__device__ __forceinline__ void mbar_expect_tx_a(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_a(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void bulk_g2s_a(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar, uint64_t policy) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(policy) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys(unsigned *p, unsigned v) { asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned ld_relaxed_sys(const unsigned *p) {
    unsigned v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ unsigned ld_relaxed(const unsigned *p) {
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }   // __threadfence() is fence.sc: dearer
__device__ __forceinline__ unsigned atom_add_relaxed(unsigned *p, unsigned v) {
    unsigned old;
    asm volatile("atom.relaxed.gpu.global.add.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
    return old;
}
__device__ __forceinline__ unsigned atom_exch_relaxed(float *p, float v) {
    unsigned old;
    asm volatile("atom.relaxed.gpu.global.exch.b32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(__float_as_uint(v)) : "memory");
    return old;
}
__device__ __forceinline__ unsigned atom_exch_relaxed_sys(float *p, float v) {
    unsigned old;
    asm volatile("atom.relaxed.sys.global.exch.b32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(__float_as_uint(v)) : "memory");
    return old;
}
__device__ __forceinline__ unsigned atom_exch64_relaxed_sys(char *p, float lo, float hi) {   // two adjacent fp32 rows in one exchange
    unsigned long long old;
    const unsigned long long v = (unsigned long long)__float_as_uint(lo) | ((unsigned long long)__float_as_uint(hi) << 32);
    asm volatile("atom.relaxed.sys.global.exch.b64 %0, [%1], %2;" : "=l"(old) : "l"(p), "l"(v) : "memory");
    return (unsigned)old ^ (unsigned)(old >> 32);
}
__device__ __forceinline__ void bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void bar_arrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }
#ifndef TB200_PLAN_CNT_STRIDE
#define TB200_PLAN_CNT_STRIDE 1
#endif
#ifndef TB200_PLAN_DEBUG
#define TB200_PLAN_DEBUG 0
#endif
constexpr bool kDbg = TB200_PLAN_DEBUG != 0;             // phase stamps (tb200_plan_trace) and the TB200_PLAN_FLAGS experiments: a probe build only
                                                          // (tools/build_variant.sh ... -DTB200_PLAN_DEBUG=1); their branches and live registers cost the step ~1-2 %
constexpr int kCS = TB200_PLAN_CNT_STRIDE;                   // words between the arrival counters of consecutive groups
constexpr int kTraceSlots = 24;                            // tb200_plan_trace: stamps 0..15, per-group sums of warp 0 in 16..21
constexpr long long kSpinLimit = 4000000000LL;            // ~2 s of SM clocks: a wedged wait gives up and raises *err

// _mm256_cvtps_epi32 + packs (ggml-quants.c:962-977), as in tb200_quantize.cu
__device__ __forceinline__ int plan_cvt_rne_sat8(float v) {
    if (!(v >= -2147483648.f && v < 2147483648.f)) return -128;
    return max(-128, min(127, __float2int_rn(v)));
}

// The group's activations: fp32 from global (L2; written by other SMs earlier in this very launch, hence ld.global.cg)
// -> optional RMSNorm*weight or silu(gate)*up -> quantize_row_q8_0 (ggml-quants.c:940-990, bit-exact: tb200_quantize.cu)
// -> the decode kernels' planar staging form: s_lo[b] = quants 0..15, s_hi[b] = quants 16..31, s_ds[b] = {d, sum q}.
// A 4-lane group owns a 32-block (8 elements per lane: two float4 loads, two shuffle steps per reduction); all loads of a
// chunk (kActR rounds of kTW*8 blocks = 16384 elements with 16 warps) are issued before the first is consumed, so a
// 14336-wide row costs one L2 round trip. Every CTA does this redundantly: 16..57 KB from L2 per SM and group.
constexpr int kActR = 4;
__device__ __forceinline__ float silu_mul(float g, float u) { return __fmul_rn(__fdiv_rn(g, __fadd_rn(1.0f, expf(-g))), u); }   // ggml_silu_f32 (ggml.c:2393) then ggml_mul
// RMSNorm scale of the group's input row, by the whole CTA: ggml_compute_forward_rms_norm_f32 (llama.cpp/ggml/src/ggml.c:11850-11896):
// sum of x*x in double, mean as float, 1/sqrtf(mean + eps)
__device__ __forceinline__ float plan_rms_scale(const PlanGroup &G, double *s_red, float *s_scale) {
    const int K = G.nb * QK;
    const float4 *x4 = reinterpret_cast<const float4 *>(G.x);
    double s = 0.0;
    for (int i0 = 0; i0 < K / 4; i0 += kTW * 32 * 2 * kActR) {
        float4 v[2 * kActR];
#pragma unroll
        for (int u = 0; u < 2 * kActR; ++u) {
            const int i = i0 + u * kTW * 32 + (int)threadIdx.x;
            v[u] = i < K / 4 ? __ldcg(x4 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < 2 * kActR; ++u)
            s += (double)__fmul_rn(v[u].x, v[u].x) + (double)__fmul_rn(v[u].y, v[u].y) + (double)__fmul_rn(v[u].z, v[u].z) + (double)__fmul_rn(v[u].w, v[u].w);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < kTW; ++w) t += s_red[w];
        const float mean = (float)(t / (double)K);
        *s_scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, G.eps)));
    }
    __syncthreads();
    return *s_scale;
}

// kNormHere: the chunk is the whole row and the prologue is RMSNorm - the sum of squares is formed from the values just loaded
// (CTA-wide, two barriers) instead of a separate pass over the row, which cost a second L2 round trip per group.
template <bool kFused, int R, bool kNormHere = false>
__device__ __forceinline__ void stage_rounds(const PlanGroup &G, int b0, int group, int j, float scale, uint4 *s_lo, uint4 *s_hi, int2 *s_ds, long long *tr,
                                             double *s_red = nullptr, float *s_scale = nullptr) {
    const int nb = G.nb, prologue = kFused ? G.prologue : TB200_PLAN_QUANT;   // !kFused: a plan without neighbours, the prologue branches fold away
    const float4 *x4 = reinterpret_cast<const float4 *>(G.x), *u4 = reinterpret_cast<const float4 *>(G.x2), *w4 = reinterpret_cast<const float4 *>(G.nw);
    float y[R][8], t[R][8];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int b = b0 + r * kTW * 8 + group;
        float4 va = make_float4(0.f, 0.f, 0.f, 0.f), vb = va, ta = va, tb = va;
        if (b < nb) {
            const int i = b * 8 + 2 * j;
            va = __ldcg(x4 + i), vb = __ldcg(x4 + i + 1);
            if (prologue == TB200_PLAN_RMSNORM) ta = __ldg(w4 + i), tb = __ldg(w4 + i + 1);
            else if (prologue == TB200_PLAN_SWIGLU) ta = __ldcg(u4 + i), tb = __ldcg(u4 + i + 1);
        }
        y[r][0] = va.x, y[r][1] = va.y, y[r][2] = va.z, y[r][3] = va.w, y[r][4] = vb.x, y[r][5] = vb.y, y[r][6] = vb.z, y[r][7] = vb.w;
        t[r][0] = ta.x, t[r][1] = ta.y, t[r][2] = ta.z, t[r][3] = ta.w, t[r][4] = tb.x, t[r][5] = tb.y, t[r][6] = tb.z, t[r][7] = tb.w;
    }
    if (tr && b0 == 0) {                                     // trace: the first round's loads have landed
        asm volatile("" ::"f"(y[0][0]), "f"(y[0][4]));
        tr[12] = clock64();
    }
    if (kNormHere) {
        // ggml_compute_forward_rms_norm_f32 (ggml.c:11850-11896): sum of x*x in double (blocks past the row were loaded as zeros), mean as float
        double sq = 0.0;
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int e = 0; e < 8; ++e) sq += (double)__fmul_rn(y[r][e], y[r][e]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
        if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = sq;
        __syncthreads();
        if (threadIdx.x == 0) {
            double tot = 0.0;
            for (int w = 0; w < kTW; ++w) tot += s_red[w];
            const float mean = (float)(tot / (double)(nb * QK));
            *s_scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, G.eps)));
        }
        __syncthreads();
        scale = *s_scale;
    }
    if (prologue == TB200_PLAN_RMSNORM) {                    // y = x*scale (ggml_vec_scale_f32), then ggml_mul by the norm weight
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int e = 0; e < 8; ++e) y[r][e] = __fmul_rn(__fmul_rn(y[r][e], scale), t[r][e]);
    } else if (prologue == TB200_PLAN_SWIGLU) {
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int e = 0; e < 8; ++e) y[r][e] = silu_mul(y[r][e], t[r][e]);
    }
    float amax[R], d[R], id[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        float a = 0.f;
#pragma unroll
        for (int e = 0; e < 8; ++e) a = fmaxf(a, fabsf(y[r][e]));
        amax[r] = a;
    }
#pragma unroll
    for (int r = 0; r < R; ++r) amax[r] = fmaxf(amax[r], __shfl_xor_sync(0xffffffffu, amax[r], 1));
#pragma unroll
    for (int r = 0; r < R; ++r) amax[r] = fmaxf(amax[r], __shfl_xor_sync(0xffffffffu, amax[r], 2));
    // d = amax/127 is only needed by the lane that stores it (j == 0), id = 127/amax by all four: lane 0 of the group divides
    // amax by 127, the others 127 by amax - ONE IEEE division per thread instead of two - and lane 1's quotient is broadcast
#pragma unroll
    for (int r = 0; r < R; ++r) d[r] = __fdiv_rn(j == 0 ? amax[r] : 127.f, j == 0 ? 127.f : amax[r]);
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const float q = __shfl_sync(0xffffffffu, d[r], ((int)threadIdx.x & 28) | 1);
        id[r] = (amax[r] != 0.0f) ? q : 0.0f;
    }
    uint2 pk[R];
    int sum[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        if (amax[r] >= 1e-30f && amax[r] <= 3e38f) {         // |x*id| <= 127*(1+2^-23): plain round-to-nearest-even, no saturation corner.
            // v + 1.5*2^23 rounds v to an integer exactly like cvt.rni (the sum's ulp is 1) and leaves it, two's complement, in
            // the low mantissa bits: byte 0 is the int8 quant. Full-rate FADD + PRMT instead of quarter-rate F2I + mask/shift/or.
            uint32_t m[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) m[e] = __float_as_uint(__fadd_rn(__fmul_rn(y[r][e], id[r]), 12582912.f));
            pk[r].x = __byte_perm(__byte_perm(m[0], m[1], 0x0040), __byte_perm(m[2], m[3], 0x0040), 0x5410);
            pk[r].y = __byte_perm(__byte_perm(m[4], m[5], 0x0040), __byte_perm(m[6], m[7], 0x0040), 0x5410);
        } else {                                             // id overflowed or inputs are not finite: the reference's cvtps/packs saturation
            int q[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) q[e] = plan_cvt_rne_sat8(__fmul_rn(y[r][e], id[r]));
            pk[r].x = (uint32_t)(q[0] & 0xff) | ((uint32_t)(q[1] & 0xff) << 8) | ((uint32_t)(q[2] & 0xff) << 16) | ((uint32_t)q[3] << 24);
            pk[r].y = (uint32_t)(q[4] & 0xff) | ((uint32_t)(q[5] & 0xff) << 8) | ((uint32_t)(q[6] & 0xff) << 16) | ((uint32_t)q[7] << 24);
        }
        sum[r] = __dp4a((int)pk[r].x, 0x01010101, __dp4a((int)pk[r].y, 0x01010101, 0));
    }
#pragma unroll
    for (int r = 0; r < R; ++r) sum[r] += __shfl_xor_sync(0xffffffffu, sum[r], 1);
#pragma unroll
    for (int r = 0; r < R; ++r) sum[r] += __shfl_xor_sync(0xffffffffu, sum[r], 2);
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int b = b0 + r * kTW * 8 + group;
        if (b < nb) {
            reinterpret_cast<uint2 *>(j < 2 ? s_lo + b : s_hi + b)[j & 1] = pk[r];
            if (j == 0) s_ds[b] = make_int2(__float_as_int(__half2float(__float2half_rn(d[r]))), sum[r]);
        }
    }
}

template <bool kFused>
__device__ __forceinline__ void stage_activations(const PlanGroup &G, uint4 *s_lo, uint4 *s_hi, int2 *s_ds, double *s_red, float *s_scale, long long *tr = nullptr) {
    const int nb = G.nb;
    const int group = threadIdx.x >> 2, j = threadIdx.x & 3;
    const int prologue = kFused ? G.prologue : TB200_PLAN_QUANT;
    const int all_rounds = (nb + kTW * 8 - 1) / (kTW * 8);
    if (prologue == TB200_PLAN_RMSNORM && all_rounds <= 2) {   // the row (K <= 8192) is one chunk: norm scale from the same registers
        if (all_rounds == 2) stage_rounds<kFused, 2, true>(G, 0, group, j, 1.f, s_lo, s_hi, s_ds, tr, s_red, s_scale);
        else stage_rounds<kFused, 1, true>(G, 0, group, j, 1.f, s_lo, s_hi, s_ds, tr, s_red, s_scale);
        return;
    }
    const float scale = prologue == TB200_PLAN_RMSNORM ? plan_rms_scale(G, s_red, s_scale) : 1.f;
    // K in chunks of up to kActR rounds (a round = kTW*8 blocks, one per 4-lane group). The rounds of a chunk are independent
    // dependency chains (loads -> amax -> two divisions -> round -> pack -> sum), so they are written phase by phase across
    // the rounds: the scheduler interleaves them and a 14336-wide row costs little more than one round's latency instead of
    // 3.5 of them (3400 -> see DESIGN 5.2c). Chunks are sized to the rounds that remain, so K = 4096 still runs one round.
    for (int b0 = 0; b0 < nb;) {
        const int rounds = (nb - b0 + kTW * 8 - 1) / (kTW * 8);
        if (rounds >= 4) stage_rounds<kFused, 4>(G, b0, group, j, scale, s_lo, s_hi, s_ds, tr);
        else if (rounds == 3) stage_rounds<kFused, 3>(G, b0, group, j, scale, s_lo, s_hi, s_ds, tr);
        else if (rounds == 2) stage_rounds<kFused, 2>(G, b0, group, j, scale, s_lo, s_hi, s_ds, tr);
        else stage_rounds<kFused, 1>(G, b0, group, j, scale, s_lo, s_hi, s_ds, tr);
        b0 += (rounds >= 4 ? 4 : rounds) * kTW * 8;
    }
}

// Block dots of one ring stage (kR rows x `segb` blocks of K) against the group's activations. kCached: the row has one
// stage (K <= 4096), so a lane meets the same four blocks in every unit and keeps their activations in registers for the
// whole group; shared memory then only carries the weights (2.3 KB per row instead of 2.3 + 5/kR).
struct ActCache { uint4 al[kSegBlocks / 32], ah[kSegBlocks / 32]; int2 ds[kSegBlocks / 32]; };
template <bool kCached, bool kFull>
__device__ __forceinline__ void stage_dots(const uint8_t *stage, const ActCache &C, const uint4 *s_lo, const uint4 *s_hi, const int2 *s_ds, int kb0, int segb,
                                           int lane, float (&acc)[kR]) {
#pragma unroll
    for (int u = 0; u < kSegBlocks / 32; ++u) {
        const int b = lane + 32 * u;
        if (kFull || b < segb) {
            uint4 al, ah;
            int2 ds;
            if (kCached) al = C.al[u], ah = C.ah[u], ds = C.ds[u];
            else al = s_lo[kb0 + b], ah = s_hi[kb0 + b], ds = s_ds[kb0 + b];
            const float dy = __int_as_float(ds.x);
            const int bias = -8 * ds.y;                      // sum (q-8)*y = sum q*y - 8*sum y (exact in int32)
#pragma unroll
            for (int r = 0; r < kR; ++r) {
                const uint4 w = reinterpret_cast<const uint4 *>(stage + r * kRowQ)[b];
                const float dw = __half2float(reinterpret_cast<const __half *>(stage + kR * kRowQ + r * kRowS)[b]);
                int l = dp4a_us(w.x & 0x0f0f0f0fu, (int)al.x, bias);
                l = dp4a_us(w.y & 0x0f0f0f0fu, (int)al.y, l);
                l = dp4a_us(w.z & 0x0f0f0f0fu, (int)al.z, l);
                l = dp4a_us(w.w & 0x0f0f0f0fu, (int)al.w, l);
                // high nibbles stay in place (= 16*q as unsigned bytes); their sum is an exact multiple of 16
                int h = dp4a_us(w.x & 0xf0f0f0f0u, (int)ah.x, 0);
                h = dp4a_us(w.y & 0xf0f0f0f0u, (int)ah.y, h);
                h = dp4a_us(w.z & 0xf0f0f0f0u, (int)ah.z, h);
                h = dp4a_us(w.w & 0xf0f0f0f0u, (int)ah.w, h);
                acc[r] = fmaf(dw * dy, (float)(l + (h >> 4)), acc[r]);
            }
        }
    }
}

// ---- Q4_K plans: Q8_K activations -----------------------------------------------------------------------------------
// nearest_int (ggml-quants.c:1632-1637) with the x86 result for a non-finite argument, as in tb200_quantize.cu
__device__ __forceinline__ int plan_nearest_int(float fval) {
    if (!(fabsf(fval) <= 3.4028234664e38f)) return 0;
    return (__float_as_int(__fadd_rn(fval, 12582912.f)) & 0x007fffff) - 0x00400000;
}
// quantize_row_q8_K (ggml-quants.c:3593-3645; bit-exact, same arithmetic as quantize_q8_K_kernel in tb200_quantize.cu) of the
// group's input after its prologue, one warp per 256-element super-block (lane l: elements 4l..4l+3 and 128+4l..128+4l+3),
// into the Q4_K decode kernel's staging form: s_q[sb][chunk c][5 x 16 B] = the 64 quants of chunk c (+ 16 B of padding: an
// 80-byte lane stride is bank-conflict free, 64 is not), s_ds[sb*8 + k] = {d, sum of the 32 quants of sub-block k}.
// R super-blocks per warp at a time (sb, sb + kTW, ...), phase by phase so that their dependency chains - loads, a five-step arg-max
// butterfly, two divisions, rounding, sums - interleave: the 70B down projection (K = 28672, 7 super-blocks per warp) would
// otherwise pay seven such chains back to back on every group boundary.
template <bool kFused, int R>
__device__ __forceinline__ void stage_q8k_rounds(const PlanGroup &G, int sb0, float scale, uint4 *s_q, int2 *s_ds) {
    const int nsb = G.nb / 8, prologue = kFused ? G.prologue : TB200_PLAN_QUANT;
    const int lane = threadIdx.x & 31;
    float v[R][8];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int sb = min(sb0 + r * kTW, nsb - 1);          // (clamped: a round past the row recomputes the last block and does not store)
        const float4 *x4 = reinterpret_cast<const float4 *>(G.x) + sb * 64;
        const float4 a = __ldcg(x4 + lane), c = __ldcg(x4 + 32 + lane);
        v[r][0] = a.x, v[r][1] = a.y, v[r][2] = a.z, v[r][3] = a.w, v[r][4] = c.x, v[r][5] = c.y, v[r][6] = c.z, v[r][7] = c.w;
        if (prologue == TB200_PLAN_RMSNORM) {
            const float4 *w4 = reinterpret_cast<const float4 *>(G.nw) + sb * 64;
            const float4 ta = __ldg(w4 + lane), tc = __ldg(w4 + 32 + lane);
            const float t[8] = {ta.x, ta.y, ta.z, ta.w, tc.x, tc.y, tc.z, tc.w};
#pragma unroll
            for (int e = 0; e < 8; ++e) v[r][e] = __fmul_rn(__fmul_rn(v[r][e], scale), t[e]);
        } else if (prologue == TB200_PLAN_SWIGLU) {
            const float4 *u4 = reinterpret_cast<const float4 *>(G.x2) + sb * 64;
            const float4 ta = __ldcg(u4 + lane), tc = __ldcg(u4 + 32 + lane);
            const float t[8] = {ta.x, ta.y, ta.z, ta.w, tc.x, tc.y, tc.z, tc.w};
#pragma unroll
            for (int e = 0; e < 8; ++e) v[r][e] = silu_mul(v[r][e], t[e]);
        }
    }
    // first arg-max |x| (strict >), carrying the signed value
    float amax[R], vmax[R];
    int imax[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        amax[r] = 0.f, vmax[r] = 0.f, imax[r] = QK_K;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float ax = fabsf(v[r][i]);
            const int idx = (i < 4) ? 4 * lane + i : 128 + 4 * lane + (i - 4);
            if (ax > amax[r]) amax[r] = ax, vmax[r] = v[r][i], imax[r] = idx;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const float oa = __shfl_xor_sync(0xffffffffu, amax[r], o), ov = __shfl_xor_sync(0xffffffffu, vmax[r], o);
            const int oi = __shfl_xor_sync(0xffffffffu, imax[r], o);
            if (oa > amax[r] || (oa == amax[r] && oi < imax[r])) amax[r] = oa, vmax[r] = ov, imax[r] = oi;
        }
    }
    float d[R], iscale[R];
#pragma unroll
    for (int r = 0; r < R; ++r) iscale[r] = amax[r] != 0.f ? __fdiv_rn(-127.f, vmax[r]) : 0.f;
#pragma unroll
    for (int r = 0; r < R; ++r) d[r] = amax[r] != 0.f ? __fdiv_rn(1.f, iscale[r]) : 0.f;
    uint32_t lo[R], hi[R];
    int s_lo[R], s_hi[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        int qv[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) qv[i] = amax[r] != 0.f ? min(127, plan_nearest_int(__fmul_rn(iscale[r], v[r][i]))) : 0;
        lo[r] = (uint32_t)(qv[0] & 0xff) | ((uint32_t)(qv[1] & 0xff) << 8) | ((uint32_t)(qv[2] & 0xff) << 16) | ((uint32_t)(qv[3] & 0xff) << 24);
        hi[r] = (uint32_t)(qv[4] & 0xff) | ((uint32_t)(qv[5] & 0xff) << 8) | ((uint32_t)(qv[6] & 0xff) << 16) | ((uint32_t)(qv[7] & 0xff) << 24);
        s_lo[r] = qv[0] + qv[1] + qv[2] + qv[3], s_hi[r] = qv[4] + qv[5] + qv[6] + qv[7];
    }
#pragma unroll
    for (int o = 1; o <= 4; o <<= 1)
#pragma unroll
        for (int r = 0; r < R; ++r) s_lo[r] += __shfl_xor_sync(0xffffffffu, s_lo[r], o), s_hi[r] += __shfl_xor_sync(0xffffffffu, s_hi[r], o);   // 32-element sums
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int sb = sb0 + r * kTW;
        if (sb < nsb) {                                      // warp-uniform
            // word `lane` of the block's 64 (elements 4*lane..): chunk lane/16, piece (lane%16)/4, word lane%4; `hi` two chunks on
            uint32_t *qw = reinterpret_cast<uint32_t *>(s_q + (size_t)sb * 20);
            const int off = ((lane >> 4) * 5 + ((lane & 15) >> 2)) * 4 + (lane & 3);
            qw[off] = lo[r], qw[off + 2 * 20] = hi[r];
            if ((lane & 7) == 0) {
                s_ds[sb * 8 + (lane >> 3)] = make_int2(__float_as_int(d[r]), s_lo[r]);
                s_ds[sb * 8 + 4 + (lane >> 3)] = make_int2(__float_as_int(d[r]), s_hi[r]);
            }
        }
    }
}
template <bool kFused>
__device__ __forceinline__ void stage_activations_q8k(const PlanGroup &G, uint4 *s_q, int2 *s_ds, double *s_red, float *s_scale) {
    const int nsb = G.nb / 8;
    const int warp = threadIdx.x >> 5;
    const float scale = (kFused && G.prologue == TB200_PLAN_RMSNORM) ? plan_rms_scale(G, s_red, s_scale) : 1.f;
    for (int sb0 = warp; sb0 < nsb;) {                       // warp-uniform
        const int rounds = (nsb - sb0 + kTW - 1) / kTW;
        if (rounds >= 4) { stage_q8k_rounds<kFused, 4>(G, sb0, scale, s_q, s_ds); sb0 += 4 * kTW; }
        else if (rounds >= 2) { stage_q8k_rounds<kFused, 2>(G, sb0, scale, s_q, s_ds); sb0 += 2 * kTW; }
        else { stage_q8k_rounds<kFused, 1>(G, sb0, scale, s_q, s_ds); sb0 += kTW; }
    }
}

// Block dots of one ring stage of Q4_K rows: kR rows x up to 16 super-blocks. A lane owns 64-element chunk c = lane % 4 of
// super-block lane / 4 (+ 8): 32 contiguous bytes of nibbles per row, the arithmetic of gemv_q4_K_kernel (tb200_gemv.cu):
// ggml_vec_dot_q4_K_q8_K (ggml-quants.c:7341-7397) - integer sub-block sums scaled by the 6-bit scales, minus dmin * sum m_j * bsum_j.
template <bool kFull>
__device__ __forceinline__ void stage_dots_q4k(const uint8_t *stage, const uint4 *s_q, const int2 *s_ds, int kb0, int segb, int lane, float (&acc)[kR]) {
    const int c = lane & 3, grp = lane >> 2, shift = 16 * (c & 1);
    const int sb0 = kb0 >> 3, nsbseg = segb >> 3;
#pragma unroll
    for (int u = 0; u < kSegBlocks / 64; ++u) {
        const int sb = grp + 8 * u;
        if (kFull || sb < nsbseg) {
            const uint4 *ap = s_q + ((size_t)(sb0 + sb) * 4 + c) * 5;
            const uint4 y0 = ap[0], y1 = ap[1], y2 = ap[2], y3 = ap[3];
            const int4 dsv = *reinterpret_cast<const int4 *>(s_ds + (sb0 + sb) * 8 + 2 * c);   // {d, sum} of sub-blocks 2c, 2c+1
            const float dy = __int_as_float(dsv.x);
#pragma unroll
            for (int r = 0; r < kR; ++r) {
                const uint4 *qrow = reinterpret_cast<const uint4 *>(stage + r * kRowQ) + sb * 8 + 2 * c;
                const uint4 a = qrow[0], b = qrow[1];
                const uint4 h = reinterpret_cast<const uint4 *>(stage + kR * kRowQ + r * kRowS)[sb];   // {d, dmin, scales[12]}
                const __half2 dd = *reinterpret_cast<const __half2 *>(&h.x);
                const float d = __low2float(dd), dmin = __high2float(dd);
                // get_scale_min_k4 (ggml-quants.c:1891-1899) four bytes at a time; the lane takes its two pairs by a fixed shift
                const uint32_t scw = c < 2 ? (h.y & 0x3f3f3f3fu) : ((h.w & 0x0f0f0f0fu) | ((h.y >> 2) & 0x30303030u));
                const uint32_t mw = c < 2 ? (h.z & 0x3f3f3f3fu) : (((h.w >> 4) & 0x0f0f0f0fu) | ((h.z >> 2) & 0x30303030u));
                const int sc_lo = (scw >> shift) & 0xff, sc_hi = (scw >> (shift + 8)) & 0xff;
                const int m_lo = (mw >> shift) & 0xff, m_hi = (mw >> (shift + 8)) & 0xff;
                int sl = dp4a_us(a.x & 0x0f0f0f0fu, (int)y0.x, 0);
                sl = dp4a_us(a.y & 0x0f0f0f0fu, (int)y0.y, sl), sl = dp4a_us(a.z & 0x0f0f0f0fu, (int)y0.z, sl), sl = dp4a_us(a.w & 0x0f0f0f0fu, (int)y0.w, sl);
                sl = dp4a_us(b.x & 0x0f0f0f0fu, (int)y1.x, sl), sl = dp4a_us(b.y & 0x0f0f0f0fu, (int)y1.y, sl);
                sl = dp4a_us(b.z & 0x0f0f0f0fu, (int)y1.z, sl), sl = dp4a_us(b.w & 0x0f0f0f0fu, (int)y1.w, sl);
                int sh = dp4a_us(a.x & 0xf0f0f0f0u, (int)y2.x, 0);                 // high nibbles in place = 16*q
                sh = dp4a_us(a.y & 0xf0f0f0f0u, (int)y2.y, sh), sh = dp4a_us(a.z & 0xf0f0f0f0u, (int)y2.z, sh), sh = dp4a_us(a.w & 0xf0f0f0f0u, (int)y2.w, sh);
                sh = dp4a_us(b.x & 0xf0f0f0f0u, (int)y3.x, sh), sh = dp4a_us(b.y & 0xf0f0f0f0u, (int)y3.y, sh);
                sh = dp4a_us(b.z & 0xf0f0f0f0u, (int)y3.z, sh), sh = dp4a_us(b.w & 0xf0f0f0f0u, (int)y3.w, sh);
                const int isum = sc_lo * sl + sc_hi * (sh >> 4);
                const int msum = m_lo * dsv.y + m_hi * dsv.w;
                acc[r] = fmaf(d * dy, (float)isum, acc[r]);
                acc[r] = fmaf(-(dmin * dy), (float)msum, acc[r]);
            }
        }
    }
}

// ---- multi-GPU plans: the cross-GPU half of a group barrier (sync warp, lane 0) --------------------------------------
// After a SHARED group every rank has stored its row slice into every rank's symmetric buffer (plain remote stores over NVLink
// from the consumer warps, written as exchanges: a warp arrives only after their old values have come back, i.e. after the peers
// hold its rows - the same protocol as for the local copy, one scope further out); the rank's LAST arriving CTA then
// publishes the group's sequence number in every peer's flag buffer, and everybody waits until all peers' numbers have arrived
// in the local flag buffer. Sequence numbers only grow (epoch * shared groups of the plan + index), so one word per peer does.
__device__ __forceinline__ void peers_signal(const PlanParams &P, unsigned seq) {
    // No fence: every warp's remote exchanges had returned before its CTA arrived, and this thread has seen all arrivals, so nothing
    // the flag could overtake is still on its way.
#pragma unroll
    for (int p = 0; p < TB200_MAX_PEERS; ++p)
        if (p < P.world && p != P.rank) st_relaxed_sys(P.peer_flag[p] + kPeerFlag0 + P.rank, seq);
}
// warp-collective: lane p waits for rank p's number, all peers at once (one lane polling them in turn paid an L2 round trip per peer
// even when every flag had long been there: ~2.5 us per shared group on 8 GPUs)
__device__ __forceinline__ void peers_wait(const PlanParams &P, unsigned seq, int lane) {
    if (lane < P.world && lane != P.rank) {
        const long long t0 = clock64();
        while ((int)(ld_relaxed_sys(P.peer_flag[P.rank] + kPeerFlag0 + lane) - seq) < 0)
            if (clock64() - t0 > kSpinLimit) { *P.err = 3; break; }
    }
    __syncwarp();
}

// kFused: some group of the plan has a neighbour folded in (RMSNorm / SwiGLU prologue, residual epilogue). Plans without any - a bare
// chain of matmuls - run the instantiation in which all of that is compiled out: the extra live registers cost the bare chain 1.8 %.
template <int kType, bool kPeers, bool kFused>
__global__ void __launch_bounds__(kTW * 32, 1) plan_kernel(const __grid_constant__ PlanParams P) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *ring = smem;                                                          // [kPW][kPS][kStageBytes]
    constexpr bool kQ4K = kType == TB200_TYPE_Q4_K;
    // staged activations. Q4_0: s_lo / s_hi [nb_max] x 16 B + s_ds [nb_max] x 8 B (40 B per 32-block); Q4_K: s_lo = the Q8_K
    // quants in the decode kernel's chunked form, [nb_max / 8] x 20 x 16 B, + s_ds [nb_max] x 8 B (48 B per 32-block)
    uint4 *s_lo = reinterpret_cast<uint4 *>(smem + kPW * kPS * kStageBytes);
    uint4 *s_hi = s_lo + P.nb_max;
    int2 *s_ds = reinterpret_cast<int2 *>(s_lo + (kQ4K ? P.nb_max / 8 * 20 : 2 * P.nb_max));
    // the descriptors of this launch's groups: every warp walks them (twice: as consumer and, a few groups ahead, as its own
    // TMA producer), and a gpu-scope fence invalidates L1, so from global memory each look-up was an L2 round trip
    const PlanGroup *tab = reinterpret_cast<const PlanGroup *>(s_ds + P.nb_max) - P.first;   // tab[g], g in [first, end)
    __shared__ uint64_t s_full[kPW][kPS];
    __shared__ double s_red[kTW];
    __shared__ float s_scale;
    constexpr int kThreads = kTW * 32;
    constexpr int kBarArrived = 1, kBarRelease = 2;         // named barriers (0 is __syncthreads)

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int GW = (int)gridDim.x * kPW, gw = (int)blockIdx.x * kPW + warp;
    const int xflags = kDbg ? P.flags : 0;              // experiment switches (probe builds only)
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(P.groups + P.first);
        uint4 *dst = reinterpret_cast<uint4 *>(s_ds + P.nb_max);
        const int n16 = (P.end - P.first) * (int)(sizeof(PlanGroup) / 16);
        for (int i = threadIdx.x; i < n16; i += kThreads) dst[i] = __ldg(src + i);
    }
    if (threadIdx.x == 0) {
        for (int i = 0; i < kPW * kPS; ++i) mbar_init(&s_full[0][0] + i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // ================================================================================================ the sync warp
    // Between two groups stands a grid-wide barrier (group g's input is a function of group g-1's output). Its protocol - wait
    // until this CTA's consumers have stored their results, release fence, arrival on the group's counter in L2, poll until
    // all CTAs have arrived, let the consumers go - runs in a warp of its own, so no consumer ever waits for a fence or an
    // atomic round trip: consumers only ARRIVE (bar.arrive, non-blocking) and use the wait for the work that is not urgent
    // (refilling the ring stage they just emptied with a row of the group after next).
    if (warp == kPW) {
        // sequence numbers of this launch's shared groups continue where the last multi-GPU plan launch on this peer group stopped
        // (one counter per peer group, in device memory: CUDA-graph replays and different plans sharing the buffers stay monotonic)
        const unsigned ep_base = kPeers ? ld_relaxed(P.epoch) - (unsigned)tab[P.first].sh_before : 0u;
        int prev_real = -1;                                 // counter of the last boundary that was a real grid barrier (to be re-armed by the next one)
        for (int g = P.first; g < P.end; ++g) {
            const PlanGroup &G = tab[g];
            long long *tr = kDbg && P.trace && lane == 0 ? P.trace + ((long)g * gridDim.x + blockIdx.x) * kTraceSlots : nullptr;
            // (descriptor look-ups of the boundary before the wait, off the critical path)
            // once any wait of this launch has given up (a peer died, the ranks' plans disagree) nobody waits any more: the launch
            // drains in microseconds instead of a spin limit per group, and tb200_plan_status() reports it
            const bool dead = g > P.first && ld_relaxed(reinterpret_cast<const unsigned *>(P.err)) != 0;
            const bool shared = kPeers && g > P.first && tab[g - 1].shared_mask != 0;   // (a dead launch still signals its peers, it only stops waiting)
            const unsigned seq = (kPeers && g > P.first) ? ep_base + (unsigned)tab[g - 1].sh_before + 1u : 0u;
            // a group declared independent starts without the grid barrier (only the CTA's own warps meet, the staging buffers are
            // reused) - unless the group before it exchanged rows with the peers, whose arrival some barrier has to cover
            const bool skip = g > P.first && G.nodep && !shared;
            if (g > P.first && !(xflags & 4) && !skip) {
                bar_sync(kBarArrived, kThreads);            // every consumer warp of this CTA has stored its results of group g-1
                if (tr) tr[1] = clock64();
                if (lane == 0) {
                    if ((xflags & 9) == 8) fence_acq_rel_gpu();   // (old protocol, flag 8: plain stores + a release fence here, ~1 us per group)
                    if (tr) tr[13] = clock64();
                    const unsigned old = atom_add_relaxed(P.cnt + (g - 1) * kCS, 1u);
                    if (tr) tr[2] = clock64();
                    if (old == gridDim.x - 1) {
                        if (prev_real >= 0) P.cnt[prev_real * kCS] = 0;    // everybody has left the previous barrier: re-arm it for the next launch
                        if (shared) peers_signal(P, seq);
                    }
                    prev_real = g - 1;
                    const long long t0 = clock64();
                    // Acquire side: a relaxed (L2) poll, no fence and no L1 invalidation per iteration. What an acquire would add is
                    // there by construction: every value another SM wrote is read with ld.global.cg / from L2 (no stale L1 copy can
                    // be hit), and those loads are issued only after the CTA barrier below, which this thread joins after the poll
                    // has seen the last arrival - whose release fence made the arriving CTA's stores visible at L2 first.
                    // (Four polls in flight a quarter of a round trip apart were measured SLOWER, 0.999 vs 0.972 ms per step: 148 pollers
                    // on the line the arrivals' atomics target already compete with them.)
                    while (!dead && ld_relaxed(P.cnt + (g - 1) * kCS) < gridDim.x) {
#ifdef TB200_PLAN_POLL_SLEEP
                        __nanosleep(TB200_PLAN_POLL_SLEEP);
#endif
                        if (clock64() - t0 > kSpinLimit) { *P.err = 1; break; }
                    }
                    if (tr) tr[3] = clock64();
                }
                __syncwarp();
                if (shared && !dead) peers_wait(P, seq, lane);
            } else if (g > P.first) bar_sync(kBarArrived, kThreads);
            bar_sync(kBarRelease, kThreads);                // consumers may stage and compute group g
            if (!(xflags & 2) || g == P.first) {
                if (kQ4K) stage_activations_q8k<kFused>(G, s_lo, s_ds, s_red, &s_scale);
                else stage_activations<kFused>(G, s_lo, s_hi, s_ds, s_red, &s_scale);
            }
            __syncthreads();
        }
        // the last group: count this CTA in, and the last CTA leaves every counter at zero for the next launch
        bar_sync(kBarArrived, kThreads);
        if (kPeers) {
            // multi-GPU plan: the launch may only end when the last group's slices of every rank have arrived here (whoever runs
            // next in the stream reads them), so the tail is a full barrier; the CTA that leaves last re-arms the counters and
            // advances the plan's epoch
            const int last = P.end - 1;
            const bool shared = tab[last].shared_mask != 0;
            const unsigned seq = ep_base + (unsigned)tab[last].sh_before + 1u;
            if (lane == 0) {
                const unsigned old = atom_add_relaxed(P.cnt + last * kCS, 1u);
                if (old == gridDim.x - 1) {
                    if (prev_real >= 0) P.cnt[prev_real * kCS] = 0;
                    if (shared) peers_signal(P, seq);
                }
                if (shared) {
                    const long long t0 = clock64();
                    while (ld_relaxed(P.cnt + last * kCS) < gridDim.x)
                        if (clock64() - t0 > kSpinLimit) { *P.err = 1; break; }
                }
            }
            __syncwarp();
            if (shared) peers_wait(P, seq, lane);
            if (lane == 0 && atom_add_relaxed(P.done, 1u) == gridDim.x - 1) {   // every CTA is past its polls of cnt[last]
                P.cnt[last * kCS] = 0;
                *P.done = 0;
                *P.epoch = seq - (shared ? 0u : 1u);   // = the number of the launch's last shared group
            }
        } else if (lane == 0 && prev_real >= 0) {           // (a launch without a grid barrier used no counter)
            const int last = P.end - 1;
            fence_acq_rel_gpu();
            const unsigned old = atom_add_relaxed(P.cnt + last * kCS, 1u);
            if (old == gridDim.x - 1) {
                P.cnt[prev_real * kCS] = 0;
                P.cnt[last * kCS] = 0;
            }
        }
        return;
    }

    // ================================================================================================ consumer warps
    const uint64_t policy = policy_evict_first();
    uint8_t *my_ring = ring + (size_t)warp * kPS * kStageBytes;

    // ---- producer side of this warp: the item (group, unit, segment) to be copied next. Lanes 0..2*kR-1 stand for (plane,
    // row): lane % kR the row of the unit, lanes < kR the nibble plane, the others the scale plane. When the unit's rows are
    // one contiguous run in memory (K = 4096, no clamped tail row) the first lane of each plane copies all kR rows at once:
    // two bulk copies per stage. What a lane needs per item lives in registers, is computed right after the iterator moves
    // (off the refill path) and is refreshed from the descriptor only when the iterator crosses into another matrix or group.
    int ig = P.first, iunit = 0, iseg = 0, i_nseg = 1, i_nunits = 0, i_nb = 0, i_ub = 0, i_ue = 0, i_rows = 1;
    const char *i_base = nullptr;
    const int prow = lane % kR, plane = (lane / kR) & 1;
    const int bunit = plane ? 2 : 16;                       // bytes per block in this lane's plane
    const int dst_off = plane ? kR * kRowQ + prow * kRowS : prow * kRowQ;
    auto first_unit = [&](const PlanGroup &G) { int p = gw - G.rot; return p < 0 ? p + GW : p; };
    auto iload_mat = [&]() {
        const PlanGroup &G = tab[ig];
        int m = 0;
#pragma unroll
        for (int mm = 1; mm < kPlanMats; ++mm)
            if (iunit >= G.unit_begin[mm]) m = mm;            // unit_begin[] of unused matrices = nunits
        i_ub = G.unit_begin[m], i_ue = G.unit_begin[m + 1], i_rows = G.rows[m];
        i_base = plane ? reinterpret_cast<const char *>(G.sc[m]) : reinterpret_cast<const char *>(G.qs[m]);
    };
    auto iseek = [&]() {   // first group from ig on in which this warp owns a unit
        while (ig < P.end) {
            const PlanGroup &G = tab[ig];
            iunit = first_unit(G), iseg = 0, i_nseg = G.nseg, i_nunits = G.nunits, i_nb = G.nb;
            if (iunit < i_nunits) { iload_mat(); return; }
            ++ig;
        }
    };
    const char *nx_src = nullptr, *nx_unit = nullptr;       // this lane's source of the next item; of segment 0 of that item's unit
    long nx_ustride = 0;                                    // bytes from a unit to this warp's next unit of the same matrix (GW units on)
    uint32_t nx_bytes = 0, nx_tx = 0;
    bool nx_on = false;
    auto iprepare = [&]() {                                 // everything from the iterator (after it moved into another matrix or group)
        const long r_first = (long)kR * (iunit - i_ub);
        const bool contig = i_nb == kSegBlocks && r_first + kR <= i_rows;          // warp-uniform
        const long row = min(r_first + prow, (long)i_rows - 1);                    // tail: the last row again, stored once
        const int b0 = iseg * kSegBlocks, segb = min(kSegBlocks, i_nb - b0);
        nx_tx = (uint32_t)segb * 18u * kR;
        nx_unit = i_base + row * i_nb * bunit;
        nx_ustride = (long)kR * GW * i_nb * bunit;
        nx_src = nx_unit + b0 * bunit;
        nx_bytes = (uint32_t)(segb * bunit) * (contig ? kR : 1);
        nx_on = lane < 2 * kR && (!contig || prow == 0) && ig < P.end;
    };
    auto iadvance = [&]() {                                 // moves the iterator and prepares the item; the common steps only add to pointers
        if (++iseg < i_nseg) {                              // next K segment of the same unit (never contiguous: rows are longer than a stage)
            const int segb = min(kSegBlocks, i_nb - iseg * kSegBlocks);
            nx_src += kSegBlocks * bunit;
            nx_tx = (uint32_t)segb * 18u * kR, nx_bytes = (uint32_t)(segb * bunit);
            return;
        }
        iseg = 0, iunit += GW;
        if (iunit < i_ue && kR * (iunit - i_ub) + kR <= i_rows) {   // a whole unit of the same matrix
            nx_unit += nx_ustride, nx_src = nx_unit;
            if (i_nseg > 1) nx_tx = (uint32_t)kSegBlocks * 18u * kR, nx_bytes = (uint32_t)(kSegBlocks * bunit);
            return;
        }
        if (iunit < i_ue) { iprepare(); return; }           // the matrix's clamped tail unit
        if (iunit < i_nunits) { iload_mat(); iprepare(); return; }
        ++ig;
        iseek();
        iprepare();
    };
    const uint32_t bar_u32 = smem_u32(&s_full[warp][0]), ring_u32 = smem_u32(my_ring) + (uint32_t)dst_off;
    auto refill = [&](int st, long long *trp = nullptr) {   // stage st is empty: copy the iterator's item into it and move on
        if (ig >= P.end) return;
        // (Lane 0 issuing all copies of an item one after the other instead of one lane per copy: 1.023 vs 0.973 ms per step.)
        if (lane < 2 * kR) {
            const uint32_t bar = bar_u32 + (uint32_t)st * 8u;
            if (lane == 0) mbar_expect_tx_a(bar, nx_tx);
            __syncwarp((1u << (2 * kR)) - 1);               // the expected byte count is posted before any copy can complete
#ifdef TB200_PLAN_TRACE_ITEMS
            if (trp) trp[14] = clock64();
#endif
            if (nx_on) bulk_g2s_a(ring_u32 + (uint32_t)st * kStageBytes, nx_src, nx_bytes, bar, policy);
#ifdef TB200_PLAN_TRACE_ITEMS
            if (trp) trp[15] = clock64();
#endif
        }
        iadvance();
    };
    iseek();
    iprepare();
#pragma unroll
    for (int st = 0; st < kPS; ++st) refill(st);

    unsigned seq = 0;                                       // items consumed by this warp
    for (int g = P.first; g < P.end; ++g) {
        const PlanGroup &G = tab[g];
        long long *tr = kDbg && P.trace && threadIdx.x == 0 ? P.trace + ((long)g * gridDim.x + blockIdx.x) * kTraceSlots : nullptr;
        if (tr) tr[0] = clock64(), tr[8] = tr[11] = 0;
        bar_sync(kBarRelease, kThreads);                    // group g-1 is complete on every SM
        if (tr) tr[4] = clock64();
        if (!(xflags & 2) || g == P.first) {
            if (kQ4K) stage_activations_q8k<kFused>(G, s_lo, s_ds, s_red, &s_scale);
            else stage_activations<kFused>(G, s_lo, s_hi, s_ds, s_red, &s_scale, tr);
        }
        if (tr) tr[7] = clock64();
        __syncthreads();
        if (tr) tr[5] = clock64();

        // ---- this warp's units of the group
        const int nb = G.nb, nseg = G.nseg, nunits = G.nunits;
        const int shmask = kPeers ? G.shared_mask : 0;     // once per group: the descriptor lives in shared memory
        const bool cached = kActCache && !kQ4K && nseg == 1;
        ActCache C;
        if (cached) {
#pragma unroll
            for (int u = 0; u < kSegBlocks / 32; ++u) {
                const int b = min(lane + 32 * u, nb - 1);
                C.al[u] = s_lo[b], C.ah[u] = s_hi[b], C.ds[u] = s_ds[b];
            }
        }
        bool first_item = tr != nullptr;
#ifdef TB200_PLAN_TRACE_ITEMS   // per-item sums cost 2.6 % of the step even when tracing is off: a probe build only (tools/build_variant.sh)
        long long *const tra = tr;
#else
        long long *const tra = nullptr;
#endif
        long long a_wait = 0, a_dots = 0, a_refill = 0, a_tail = 0, a_items = 0, t_prev = tra ? clock64() : 0;   // trace: where warp 0's time goes
        unsigned sink = 0, pend = 0;
        unsigned pendr[TB200_MAX_PEERS] = {0, 0, 0, 0, 0, 0, 0, 0};   // multi-GPU plans: the remote exchanges of the previous unit
        int deferred = -1;                                  // the stage emptied by the warp's last item of the group: refilled after the arrival
        for (int unit = first_unit(G); unit < nunits; unit += GW) {
            float acc[kR];
#pragma unroll
            for (int r = 0; r < kR; ++r) acc[r] = 0.f;
            // where the unit's results go (looked up while the stage is still on its way): row r of the unit ends up in lane r*(32/kR)
            int m = 0;
#pragma unroll
            for (int mm = 1; mm < kPlanMats; ++mm)
                if (unit >= G.unit_begin[mm]) m = mm;
            const int my_r = lane / (32 / kR);
            const int o_row = kR * (unit - G.unit_begin[m]) + my_r;
            const bool o_on = lane % (32 / kR) == 0 && o_row < G.rows[m];
            float *o_ptr = G.out[m] + o_row;
            const float *o_res = (kFused && G.res[m]) ? G.res[m] + o_row : nullptr;
            const float resv = (kFused && o_on && o_res) ? __ldcg(o_res) : 0.f;   // fetched now, used after the unit's block dots: an in-order warp would otherwise park on this L2 round trip at every unit's tail
            const bool last_unit = unit + GW >= nunits;
            for (int seg = 0; seg < nseg; ++seg, ++seq) {
                const int st = seq % kPS;
                {
                    const uint32_t bar = bar_u32 + (uint32_t)st * 8u;
                    const uint32_t parity = (seq / kPS) & 1;
                    if (!mbar_try_wait_a(bar, parity)) {
                        const long long t0 = clock64();
                        while (!mbar_try_wait_a(bar, parity))
                            if (clock64() - t0 > kSpinLimit) { *P.err = 2; break; }
                    }
                }
                if (first_item) tr[8] = clock64();
                if (tra) { const long long t = clock64(); a_wait += t - t_prev, t_prev = t; }
                const uint8_t *stage = my_ring + st * kStageBytes;
                const int kb0 = seg * kSegBlocks, segb = min(kSegBlocks, nb - kb0);
                if (kQ4K) {
                    if (segb == kSegBlocks) stage_dots_q4k<true>(stage, s_lo, s_ds, kb0, segb, lane, acc);
                    else stage_dots_q4k<false>(stage, s_lo, s_ds, kb0, segb, lane, acc);
                } else if (cached) {
                    if (segb == kSegBlocks) stage_dots<true, true>(stage, C, s_lo, s_hi, s_ds, kb0, segb, lane, acc);
                    else stage_dots<true, false>(stage, C, s_lo, s_hi, s_ds, kb0, segb, lane, acc);
                } else {
                    if (segb == kSegBlocks) stage_dots<false, true>(stage, C, s_lo, s_hi, s_ds, kb0, segb, lane, acc);
                    else stage_dots<false, false>(stage, C, s_lo, s_hi, s_ds, kb0, segb, lane, acc);
                }
                __syncwarp();                               // every lane has read the stage: it can be refilled
                if (first_item) tr[9] = clock64();
                if (tra) { const long long t = clock64(); a_dots += t - t_prev, t_prev = t, ++a_items; }
                if (last_unit && seg == nseg - 1) deferred = st;   // not urgent (it is for a later group): after the arrival
                else refill(st, first_item ? tr : nullptr);
                if (first_item) tr[10] = clock64(), first_item = false;
                if (tra) { const long long t = clock64(); a_refill += t - t_prev, t_prev = t; }
            }
            // kR sums with one transposing butterfly: each step halves the values a lane carries (the lane keeps the rows of its own
            // half and hands the others over), then plain steps finish the sum inside a 32/kR-lane segment: 5 or 6 shuffles, not 5*kR
            float v;
            if (kR == 2) {
                const bool up = lane & 16;
                v = (up ? acc[1] : acc[0]) + __shfl_xor_sync(0xffffffffu, up ? acc[0] : acc[1], 16);
            } else {
                const bool up = lane & 16, up2 = lane & 8;
                const float a = (up ? acc[2] : acc[0]) + __shfl_xor_sync(0xffffffffu, up ? acc[0] : acc[2], 16);
                const float b = (up ? acc[3] : acc[1]) + __shfl_xor_sync(0xffffffffu, up ? acc[1] : acc[3], 16);
                v = (up2 ? b : a) + __shfl_xor_sync(0xffffffffu, up2 ? a : b, 8);
            }
#pragma unroll
            for (int o = 16 / kR; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            const float val = (kFused && o_on && o_res) ? __fadd_rn(v, resv) : v;
            if (o_on) {
                // The result is written with an exchange, not a store: when the old value comes back the write has been performed
                // at L2, the point of coherence of every SM. Once all of a warp's exchanges have returned it arrives; the CTA's
                // arrival on the group counter (one relaxed atomic by the sync warp, after the CTA barrier) then needs no release
                // fence - a fence.acq_rel.gpu there cost ~1 us per group, it waits for the SM's whole memory pipe including the
                // ring's bulk copies in flight. The round trips overlap the following units' block dots; only the last is exposed.
                // (A warp executes in order: consuming the old value right away would park the warp for the whole round trip,
                // ~600 cycles per unit. It is folded in one unit later, when it has long since returned.)
                if (!(xflags & 8)) sink ^= pend, pend = atom_exch_relaxed(o_ptr, val);
                else *o_ptr = val;
                if (tr && tr[11] == 0) tr[11] = clock64();
            }
            if (kPeers && ((shmask >> m) & 1)) {
                // row-sharded matrix: the same rows of every rank's symmetric buffer, over NVLink - again as exchanges whose old
                // values are folded in one unit later: once they are back the peers hold the rows, so no system-scope fence is
                // needed before the arrival (a fence.acq_rel.sys per CTA and group cost 3.8 us: 1.41 vs 0.93 ms per step with
                // every group of the 8B step shared on two GPUs; plain stores cost nothing measurable but prove nothing). Lane 0
                // sends the unit's row pair as ONE 8-byte exchange per peer (half the NVLink atomics; a lone tail row as 4 bytes).
                static_assert(kR == 2, "the pair exchange assumes two rows per unit");
                const float val1 = __shfl_sync(0xffffffffu, val, 16);
                const int row0 = kR * (unit - G.unit_begin[m]);
                const bool pair = row0 + 1 < G.rows[m] && ((reinterpret_cast<uintptr_t>(G.out[m] + row0) & 7) == 0);   // warp-uniform
                if (o_on && (pair ? lane == 0 : true)) {
                    const long off = reinterpret_cast<const char *>(o_ptr) - P.local_base;
#pragma unroll
                    for (int p = 0; p < TB200_MAX_PEERS; ++p)
                        if (p < P.world && p != P.rank) {
                            sink ^= pendr[p];
                            if (pair) pendr[p] = atom_exch64_relaxed_sys(P.peer_base[p] + off, val, val1);
                            else pendr[p] = atom_exch_relaxed_sys(reinterpret_cast<float *>(P.peer_base[p] + off), val);
                        }
                }
            }
            if (tra) { const long long t = clock64(); a_tail += t - t_prev, t_prev = t; }
        }
        if (tra) tr[16] = a_wait, tr[17] = a_dots, tr[18] = a_refill, tr[19] = a_tail, tr[20] = a_items;
        if (kPeers && shmask != 0) {
#pragma unroll
            for (int p = 0; p < TB200_MAX_PEERS; ++p)
                if (p < P.world) sink ^= pendr[p], pendr[p] = 0;
        }
        if (!(xflags & 8)) { sink ^= pend; asm volatile("" ::"r"(sink) : "memory"); }   // every exchange of this lane has returned

        if (tr) tr[6] = clock64();
        __syncwarp();
        bar_arrive(kBarArrived, kThreads);                  // this warp's results of group g are stored; never blocks
        if (deferred >= 0) refill(deferred);
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------------ host side
typedef void (*PlanKernelFn)(const PlanParams);
static PlanKernelFn plan_kernel_for(int type, bool peers, bool fused) {
    const bool k = type == TB200_TYPE_Q4_K;
    if (fused) return k ? (peers ? plan_kernel<TB200_TYPE_Q4_K, true, true> : plan_kernel<TB200_TYPE_Q4_K, false, true>)
                        : (peers ? plan_kernel<TB200_TYPE_Q4_0, true, true> : plan_kernel<TB200_TYPE_Q4_0, false, true>);
    return k ? (peers ? plan_kernel<TB200_TYPE_Q4_K, true, false> : plan_kernel<TB200_TYPE_Q4_K, false, false>)
             : (peers ? plan_kernel<TB200_TYPE_Q4_0, true, false> : plan_kernel<TB200_TYPE_Q4_0, false, false>);
}
static bool plan_kernels_set_smem() {
    bool ok = true;
    for (int t = 0; t < 2; ++t)
        for (int pe = 0; pe < 2; ++pe)
            for (int f = 0; f < 2; ++f)
                ok = ok && cudaFuncSetAttribute(plan_kernel_for(t ? TB200_TYPE_Q4_K : TB200_TYPE_Q4_0, pe != 0, f != 0), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPlanSmemMax) == cudaSuccess;
    return ok;
}

struct tb200_plan {
    int n_groups, nb_max, device, type;
    int world, rank, n_shared;
    bool fused;                        // some group has a prologue other than QUANT or a residual
    const char *local_base;
    char *peer_base[TB200_MAX_PEERS];
    unsigned *peer_flag[TB200_MAX_PEERS];
    unsigned *d_epoch;                 // multi-GPU plans: the peer group's sequence counter
    PlanGroup *d_groups;
    unsigned *d_cnt;
    int *d_err;
    long long *d_trace;
    int grid;
    size_t smem;
};

extern "C" {

static tb200_plan *plan_create_impl(const tb200_peers *pg, size_t sym_bytes, int n_groups, const tb200_plan_group *groups) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return nullptr;
    if (n_groups < 1 || !groups) { tb200_set_error("plan_create: bad arguments"); return nullptr; }
    PlanGroup *h = (PlanGroup *)calloc((size_t)n_groups, sizeof(PlanGroup));
    if (!h) return nullptr;
    const long GW = (long)c->sm_count * kPW;
    long rot = 0;
    int nb_max = 0, n_shared = 0;
    bool any_fused = false;
    const char *sym_lo = pg ? (const char *)pg->out_base[pg->rank] : nullptr, *sym_hi = sym_lo + sym_bytes;
    for (int g = 0; g < n_groups; ++g) {
        const tb200_plan_group &s = groups[g];
        PlanGroup &G = h[g];
        const char *bad = nullptr;
        if (s.n < 1 || s.n > kPlanMats || !s.x) bad = "1..4 matrices and an input are required";
        else if ((s.prologue & ~TB200_PLAN_INDEPENDENT) < TB200_PLAN_QUANT || (s.prologue & ~TB200_PLAN_INDEPENDENT) > TB200_PLAN_SWIGLU) bad = "unknown prologue";
        else if ((s.prologue & 0xff) == TB200_PLAN_RMSNORM && !s.norm_weight) bad = "RMSNORM needs norm_weight";
        else if ((s.prologue & 0xff) == TB200_PLAN_SWIGLU && !s.x2) bad = "SWIGLU needs x2";
        else if (((uintptr_t)s.x | (uintptr_t)s.x2 | (uintptr_t)s.norm_weight) & 15) bad = "inputs must be 16-byte aligned";
        long units = 0;
        for (int m = 0; m < s.n && !bad; ++m) {
            const tb200_weights *w = s.w[m];
            if (!w || (w->type != TB200_TYPE_Q4_0 && w->type != TB200_TYPE_Q4_K)) bad = "only Q4_0 or Q4_K weights (use the per-matrix API for the other types)";
            else if (w->type != groups[0].w[0]->type) bad = "all matrices of a plan have one weight type";
            else if (w->k != s.w[0]->k || w->k % 256) bad = "matrices of a group share K, a multiple of 256";
            else if (w->rows < 1 || !s.out[m]) bad = "empty matrix or missing output";
            else if (w->device != c->device) bad = "weights live on another device";
            if (bad) break;
            G.unit_begin[m] = (int)units;
            units += (w->rows + kR - 1) / kR;
            G.rows[m] = (int)w->rows, G.qs[m] = (const uint4 *)w->qs, G.sc[m] = (const __half *)w->sc, G.out[m] = s.out[m], G.res[m] = s.residual[m];
            // multi-GPU plan: an output inside this rank's symmetric buffer is a row slice that every rank must receive
            const char *o = (const char *)s.out[m];
            if (pg && pg->world > 1 && o >= sym_lo && o < sym_hi) {
                if (o + w->rows * 4 > sym_hi) bad = "output slice runs past the symmetric buffer";
                G.shared_mask |= 1 << m;
            }
        }
        if (bad) { tb200_set_error("plan_create: group %d: %s", g, bad); free(h); return nullptr; }
        for (int m = s.n; m <= kPlanMats; ++m) G.unit_begin[m] = (int)units;
        G.x = s.x, G.x2 = s.x2, G.nw = s.norm_weight, G.eps = s.eps, G.prologue = s.prologue & 0xff, G.nodep = (s.prologue & TB200_PLAN_INDEPENDENT) != 0;
        G.nmat = s.n, G.nb = (int)(s.w[0]->k / QK), G.nseg = (G.nb + kSegBlocks - 1) / kSegBlocks;
        G.sh_before = n_shared;
        if ((s.prologue & 0xff) != TB200_PLAN_QUANT) any_fused = true;
        for (int m = 0; m < s.n; ++m) any_fused = any_fused || s.residual[m] != nullptr;
        n_shared += G.shared_mask != 0;
        G.nunits = (int)units, G.rot = (int)(rot % GW);
        rot += units;                                       // the units left over by one group go to the warps after them in the next
        nb_max = G.nb > nb_max ? G.nb : nb_max;
    }
    tb200_plan *p = new tb200_plan();
    memset(p, 0, sizeof(*p));
    p->n_groups = n_groups, p->nb_max = nb_max, p->device = c->device, p->type = groups[0].w[0]->type;
    p->smem = (size_t)kPW * kPS * kStageBytes + (size_t)nb_max * (p->type == TB200_TYPE_Q4_K ? 48 : 40);   // + the launch's descriptor table (tb200_plan_run)
    p->fused = any_fused;
    p->world = pg ? pg->world : 1, p->rank = pg ? pg->rank : 0, p->n_shared = n_shared, p->local_base = sym_lo;
    for (int q = 0; pg && q < pg->world; ++q) p->peer_base[q] = (char *)pg->out_base[q], p->peer_flag[q] = pg->flags[q];
    bool ok = p->smem + 16 * sizeof(PlanGroup) <= kPlanSmemMax;
    if (!ok) tb200_set_error("plan_create: K = %d does not fit the activation staging", nb_max * QK);
    ok = ok && cudaMalloc((void **)&p->d_groups, (size_t)n_groups * sizeof(PlanGroup)) == cudaSuccess;
    ok = ok && cudaMalloc((void **)&p->d_cnt, (size_t)(n_groups + 1) * kCS * 4 + 16) == cudaSuccess;
    if (ok) {
        p->d_err = (int *)(p->d_cnt + (size_t)(n_groups + 1) * kCS);   // then {-, done, epoch}
        p->d_epoch = pg ? pg->plan_seq : p->d_cnt + (size_t)(n_groups + 1) * kCS + 3;
        ok = cudaMemcpy(p->d_groups, h, (size_t)n_groups * sizeof(PlanGroup), cudaMemcpyHostToDevice) == cudaSuccess &&
             cudaMemset(p->d_cnt, 0, (size_t)(n_groups + 1) * kCS * 4 + 16) == cudaSuccess &&
             plan_kernels_set_smem();
        if (!ok) tb200_set_error("plan_create: %s", cudaGetErrorString(cudaGetLastError()));
    } else cudaGetLastError();
    free(h);
    if (!ok) { tb200_plan_free(p); return nullptr; }
    return p;
}

tb200_plan *tb200_plan_create(int n_groups, const tb200_plan_group *groups) { return plan_create_impl(nullptr, 0, n_groups, groups); }

tb200_plan *tb200_plan_create_peers(tb200_peers *g, size_t symmetric_bytes, int n_groups, const tb200_plan_group *groups) {
    if (!g || !symmetric_bytes) { tb200_set_error("plan_create_peers: bad arguments"); return nullptr; }
    return plan_create_impl(g, symmetric_bytes, n_groups, groups);
}

int tb200_plan_run(tb200_plan *p, int first, int count) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    if (!p || first < 0 || count < 0 || first + count > p->n_groups || p->device != c->device) { tb200_set_error("plan_run: bad arguments"); return TB200_ERR_INVALID; }
    // the descriptors of a launch live in shared memory: longer sequences go out as several launches
    const int per_launch = (int)((kPlanSmemMax - p->smem) / sizeof(PlanGroup));
    for (int g0 = first; g0 < first + count; g0 += per_launch) {
        const int n = first + count - g0 < per_launch ? first + count - g0 : per_launch;
        PlanParams P;
        P.groups = p->d_groups, P.first = g0, P.end = g0 + n, P.cnt = p->d_cnt, P.err = p->d_err, P.nb_max = p->nb_max;
        P.trace = p->d_trace;
        P.world = p->world, P.rank = p->rank, P.n_shared = p->n_shared, P.local_base = p->local_base;
        for (int q = 0; q < TB200_MAX_PEERS; ++q) P.peer_base[q] = p->peer_base[q], P.peer_flag[q] = p->peer_flag[q];
        P.done = p->d_cnt + (size_t)(p->n_groups + 1) * kCS + 2, P.epoch = p->d_epoch;
        static const int flags = [] { const char *v = getenv("TB200_PLAN_FLAGS"); return v && *v ? atoi(v) : 0; }();
        P.flags = flags;
        p->grid = c->sm_count;
        // one CTA per SM, all co-resident (the groups synchronise through arrival counters); plain stream order, no PDL: the
        // kernel owns the whole GPU for the step
        const size_t smem = p->smem + (size_t)n * sizeof(PlanGroup);
        const dim3 grid((unsigned)c->sm_count), block(kTW * 32);
        static const bool force_peers = [] { const char *v = getenv("TB200_PLAN_FORCE_PEERS"); return v && *v == '1'; }();   // A/B: the multi-GPU instantiation on one GPU
        plan_kernel_for(p->type, p->world > 1 || force_peers, p->fused)<<<grid, block, smem, c->stream>>>(P);
        tb200_count_launch();
        TB_CUDA(cudaGetLastError());
    }
    return TB200_OK;
}

int tb200_plan_status(tb200_plan *p) {
    if (!p) return -1;
    int v = 0;
    if (cudaMemcpy(&v, p->d_err, 4, cudaMemcpyDeviceToHost) != cudaSuccess) { cudaGetLastError(); return -1; }
    return v;
}

int tb200_plan_reset(tb200_plan *p) {
    if (!p) return TB200_ERR_INVALID;
    TB_CUDA(cudaMemset(p->d_cnt, 0, (size_t)(p->n_groups + 1) * kCS * 4 + 12));   // counters, the error word and `done`; the epoch of a multi-GPU plan keeps counting (the peers' flags do)
    return TB200_OK;
}

// Debug aid: per group and CTA, the SM clock at {loop top, CTA done with g-1, arrived, barrier complete, CTA released,
// activations staged, warp 0 done}. Call once before tb200_plan_run to switch it on; copies out after a run. out: [n_groups][grid][24].
long tb200_plan_trace(tb200_plan *p, long long *out, long capacity) {
    tb200_ctx *c = tb200_get_ctx();
    if (!p || !c) return -1;
    if (!kDbg) { tb200_set_error("plan_trace: this build has no phase stamps (compile tb200_plan.cu with -DTB200_PLAN_DEBUG=1)"); return -2; }
    const long n = (long)p->n_groups * c->sm_count * kTraceSlots;
    if (!p->d_trace) {
        if (cudaMalloc((void **)&p->d_trace, (size_t)n * 8) != cudaSuccess || cudaMemset(p->d_trace, 0, (size_t)n * 8) != cudaSuccess) { cudaGetLastError(); p->d_trace = nullptr; return -1; }
        return 0;
    }
    if (!out || capacity < n) return n;
    if (cudaMemcpy(out, p->d_trace, (size_t)n * 8, cudaMemcpyDeviceToHost) != cudaSuccess) { cudaGetLastError(); return -1; }
    return n;
}

void tb200_plan_free(tb200_plan *p) {
    if (!p) return;
    if (p->d_trace) cudaFree(p->d_trace);
    if (p->d_groups) cudaFree(p->d_groups);
    if (p->d_cnt) cudaFree(p->d_cnt);
    delete p;
}

}  // extern "C"
