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

constexpr int kPW = 16;                                   // warps per CTA, every one a consumer with its own ring
constexpr int kPS = 2;                                    // ring stages per warp
constexpr int kSegBlocks = 128;                           // 32-element blocks of K per stage and row (4096 elements)
constexpr int kRowQ = kSegBlocks * 16, kRowS = kSegBlocks * 2;
constexpr int kStageBytes = 2 * (kRowQ + kRowS);          // a row PAIR: 4608 B
constexpr int kPlanMats = TB200_GEMV_MAX_MATS;

struct PlanGroup {
    const float *x, *x2, *nw;          // prologue inputs (see tb200_plan_group)
    float eps;
    int prologue;
    int nb, nseg;                      // 32-blocks per row; ring stages per row pair
    int npairs, rot;                   // row pairs of all matrices; rotation of the warp -> pair assignment
    int nmat;
    int pair_begin[kPlanMats + 1];
    int rows[kPlanMats];
    const uint4 *qs[kPlanMats];
    const __half *sc[kPlanMats];
    float *out[kPlanMats];
    const float *res[kPlanMats];
};

struct PlanParams {
    const PlanGroup *groups;
    int first, end;                    // groups [first, end) run in this launch
    unsigned *cnt;                     // [n_groups] arrival counters, all zero between launches
    int *err;                          // set when a wait gave up (the results of that launch are then undefined)
    int nb_max;
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
// 1-D bulk copy global -> shared, completion counted in bytes on an mbarrier; the weights are read exactly once, so the
// lines are marked evict-first in L2 and leave the activations / outputs of the step resident there
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar, uint64_t policy) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy) : "memory");
}
__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ unsigned ld_acquire(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
constexpr long long kSpinLimit = 4000000000LL;            // ~2 s of SM clocks: a wedged wait gives up and raises *err

// _mm256_cvtps_epi32 + packs (ggml-quants.c:962-977), as in tb200_quantize.cu
__device__ __forceinline__ int plan_cvt_rne_sat8(float v) {
    if (!(v >= -2147483648.f && v < 2147483648.f)) return -128;
    return max(-128, min(127, __float2int_rn(v)));
}

// The group's activations: fp32 from global (L2; written by other SMs earlier in this very launch, hence ld.global.cg)
// -> optional RMSNorm*weight or silu(gate)*up -> quantize_row_q8_0 (ggml-quants.c:940-990, bit-exact: tb200_quantize.cu)
// -> the decode kernels' planar staging form: s_lo[b] = quants 0..15, s_hi[b] = quants 16..31, s_ds[b] = {d, sum q}.
__device__ __forceinline__ void stage_activations(const PlanGroup &G, uint4 *s_lo, uint4 *s_hi, int2 *s_ds, double *s_red, float *s_scale) {
    const int nb = G.nb, K = nb * QK;
    float scale = 1.f;
    if (G.prologue == TB200_PLAN_RMSNORM) {
        // ggml_compute_forward_rms_norm_f32 (llama.cpp/ggml/src/ggml.c): sum of x*x in double, mean as float, 1/sqrtf(mean + eps)
        double s = 0.0;
        for (int i = threadIdx.x; i < K / 4; i += kPW * 32) {
            const float4 v = __ldcg(reinterpret_cast<const float4 *>(G.x) + i);
            s += (double)__fmul_rn(v.x, v.x) + (double)__fmul_rn(v.y, v.y) + (double)__fmul_rn(v.z, v.z) + (double)__fmul_rn(v.w, v.w);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = s;
        __syncthreads();
        if (threadIdx.x == 0) {
            double t = 0.0;
            for (int w = 0; w < kPW; ++w) t += s_red[w];
            const float mean = (float)(t / (double)K);
            *s_scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, G.eps)));
        }
        __syncthreads();
        scale = *s_scale;
    }
    const int group = threadIdx.x >> 3, g8 = threadIdx.x & 7;
    for (int b0 = 0; b0 < nb; b0 += kPW * 4) {
        const int b = b0 + group;
        const bool live = b < nb;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (live) {
            const int i4 = b * 8 + g8;
            v = __ldcg(reinterpret_cast<const float4 *>(G.x) + i4);
            if (G.prologue == TB200_PLAN_RMSNORM) {   // y = x*scale (ggml_vec_scale_f32), then ggml_mul by the norm weight
                const float4 w = __ldg(reinterpret_cast<const float4 *>(G.nw) + i4);
                v.x = __fmul_rn(__fmul_rn(v.x, scale), w.x), v.y = __fmul_rn(__fmul_rn(v.y, scale), w.y);
                v.z = __fmul_rn(__fmul_rn(v.z, scale), w.z), v.w = __fmul_rn(__fmul_rn(v.w, scale), w.w);
            } else if (G.prologue == TB200_PLAN_SWIGLU) {   // ggml_silu(gate) then ggml_mul(., up): x/(1+exp(-x)) * up
                const float4 u = __ldcg(reinterpret_cast<const float4 *>(G.x2) + i4);
                v.x = __fmul_rn(__fdiv_rn(v.x, __fadd_rn(1.0f, expf(-v.x))), u.x), v.y = __fmul_rn(__fdiv_rn(v.y, __fadd_rn(1.0f, expf(-v.y))), u.y);
                v.z = __fmul_rn(__fdiv_rn(v.z, __fadd_rn(1.0f, expf(-v.z))), u.z), v.w = __fmul_rn(__fdiv_rn(v.w, __fadd_rn(1.0f, expf(-v.w))), u.w);
            }
        }
        float amax = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w)));
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
        const float d = __fdiv_rn(amax, 127.f);
        const float id = (amax != 0.0f) ? __fdiv_rn(127.f, amax) : 0.0f;
        const int q0 = plan_cvt_rne_sat8(__fmul_rn(v.x, id)), q1 = plan_cvt_rne_sat8(__fmul_rn(v.y, id));
        const int q2 = plan_cvt_rne_sat8(__fmul_rn(v.z, id)), q3 = plan_cvt_rne_sat8(__fmul_rn(v.w, id));
        const uint32_t packed = (uint32_t)(q0 & 0xff) | ((uint32_t)(q1 & 0xff) << 8) | ((uint32_t)(q2 & 0xff) << 16) | ((uint32_t)(q3 & 0xff) << 24);
        int sum = q0 + q1 + q2 + q3;
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        if (live) {
            reinterpret_cast<uint32_t *>(g8 < 4 ? s_lo + b : s_hi + b)[g8 & 3] = packed;
            if (g8 == 0) s_ds[b] = make_int2(__float_as_int(__half2float(__float2half_rn(d))), sum);
        }
    }
}

__global__ void __launch_bounds__(kPW * 32, 1) plan_q4_0_kernel(const __grid_constant__ PlanParams P) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *ring = smem;                                                          // [kPW][kPS][kStageBytes]
    uint4 *s_lo = reinterpret_cast<uint4 *>(smem + kPW * kPS * kStageBytes);       // [nb_max]
    uint4 *s_hi = s_lo + P.nb_max;
    int2 *s_ds = reinterpret_cast<int2 *>(s_hi + P.nb_max);
    __shared__ uint64_t s_full[kPW][kPS];
    __shared__ double s_red[kPW];
    __shared__ float s_scale;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int GW = (int)gridDim.x * kPW, gw = (int)blockIdx.x * kPW + warp;
    if (threadIdx.x == 0) {
        for (int i = 0; i < kPW * kPS; ++i) mbar_init(&s_full[0][0] + i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const uint64_t policy = policy_evict_first();
    uint8_t *my_ring = ring + (size_t)warp * kPS * kStageBytes;

    // ---- producer side of this warp: the item (group, pair, segment) to be copied next
    int ig = P.first, ipair = 0, iseg = 0, i_nseg = 1, i_npairs = 0;
    auto first_pair = [&](const PlanGroup &G) { int p = gw - G.rot % GW; return p < 0 ? p + GW : p; };
    auto iseek = [&]() {   // first group from ig on in which this warp owns a pair
        while (ig < P.end) {
            const PlanGroup &G = P.groups[ig];
            ipair = first_pair(G), iseg = 0, i_nseg = G.nseg, i_npairs = G.npairs;
            if (ipair < i_npairs) return;
            ++ig;
        }
    };
    auto issue = [&](int st) {   // lane 0: four bulk copies (two rows x {nibbles, scales}) into stage st
        const PlanGroup &G = P.groups[ig];
        int m = 0;
#pragma unroll
        for (int mm = 1; mm < kPlanMats; ++mm)
            if (mm < G.nmat && ipair >= G.pair_begin[mm]) m = mm;
        const int lp = ipair - G.pair_begin[m];
        const long r0 = 2L * lp, r1 = min(2L * lp + 1, (long)G.rows[m] - 1);   // odd tail: the last row twice, stored once
        const int b0 = iseg * kSegBlocks, segb = min(kSegBlocks, G.nb - b0);
        uint8_t *stage = my_ring + st * kStageBytes;
        uint64_t *bar = &s_full[warp][st];
        mbar_expect_tx(bar, (uint32_t)segb * 36u);
        bulk_g2s(stage, G.qs[m] + (r0 * G.nb + b0), segb * 16, bar, policy);
        bulk_g2s(stage + kRowQ, G.qs[m] + (r1 * G.nb + b0), segb * 16, bar, policy);
        bulk_g2s(stage + 2 * kRowQ, G.sc[m] + (r0 * G.nb + b0), segb * 2, bar, policy);
        bulk_g2s(stage + 2 * kRowQ + kRowS, G.sc[m] + (r1 * G.nb + b0), segb * 2, bar, policy);
    };
    auto iadvance = [&]() {
        if (++iseg < i_nseg) return;
        iseg = 0, ipair += GW;
        if (ipair < i_npairs) return;
        ++ig;
        iseek();
    };
    iseek();
#pragma unroll
    for (int st = 0; st < kPS; ++st)
        if (ig < P.end) {
            if (lane == 0) issue(st);
            iadvance();
        }

    unsigned seq = 0;                                       // items consumed by this warp
    for (int g = P.first; g < P.end; ++g) {
        const PlanGroup &G = P.groups[g];
        // ---- the previous group is complete on every SM (its outputs are this group's inputs)
        __syncthreads();                                    // this CTA's warps are done with group g-1 and its staged activations
        if (g > P.first && threadIdx.x == 0) {
            __threadfence();
            const unsigned old = atomicAdd(P.cnt + (g - 1), 1u);
            if (old == gridDim.x - 1 && g - 1 > P.first) P.cnt[g - 2] = 0;   // everybody has left barrier g-2: re-arm it for the next launch
            const long long t0 = clock64();
            while (ld_acquire(P.cnt + (g - 1)) < gridDim.x)
                if (clock64() - t0 > kSpinLimit) { *P.err = 1; break; }
        }
        __syncthreads();
        stage_activations(G, s_lo, s_hi, s_ds, s_red, &s_scale);
        __syncthreads();

        // ---- this warp's row pairs of the group
        const int nb = G.nb, nseg = G.nseg, npairs = G.npairs;
        for (int pair = first_pair(G); pair < npairs; pair += GW) {
            float acc0 = 0.f, acc1 = 0.f;
            for (int seg = 0; seg < nseg; ++seg, ++seq) {
                const int st = seq % kPS;
                {
                    uint64_t *bar = &s_full[warp][st];
                    const uint32_t parity = (seq / kPS) & 1;
                    if (!mbar_try_wait(bar, parity)) {
                        const long long t0 = clock64();
                        while (!mbar_try_wait(bar, parity))
                            if (clock64() - t0 > kSpinLimit) { *P.err = 2; break; }
                    }
                }
                const uint8_t *stage = my_ring + st * kStageBytes;
                const uint4 *q0 = reinterpret_cast<const uint4 *>(stage), *q1 = reinterpret_cast<const uint4 *>(stage + kRowQ);
                const __half *d0 = reinterpret_cast<const __half *>(stage + 2 * kRowQ), *d1 = reinterpret_cast<const __half *>(stage + 2 * kRowQ + kRowS);
                const int kb0 = seg * kSegBlocks, segb = min(kSegBlocks, nb - kb0);
#pragma unroll
                for (int u = 0; u < kSegBlocks / 32; ++u) {
                    const int b = lane + 32 * u;
                    if (b < segb) {
                        const uint4 w0 = q0[b], w1 = q1[b];
                        const float dw0 = __half2float(d0[b]), dw1 = __half2float(d1[b]);
                        const uint4 al = s_lo[kb0 + b], ah = s_hi[kb0 + b];
                        const int2 ds = s_ds[kb0 + b];
                        const float dy = __int_as_float(ds.x);
                        const int bias = -8 * ds.y;                      // sum (q-8)*y = sum q*y - 8*sum y (exact in int32)
                        int l0 = dp4a_us(w0.x & 0x0f0f0f0fu, (int)al.x, bias), l1 = dp4a_us(w1.x & 0x0f0f0f0fu, (int)al.x, bias);
                        l0 = dp4a_us(w0.y & 0x0f0f0f0fu, (int)al.y, l0), l1 = dp4a_us(w1.y & 0x0f0f0f0fu, (int)al.y, l1);
                        l0 = dp4a_us(w0.z & 0x0f0f0f0fu, (int)al.z, l0), l1 = dp4a_us(w1.z & 0x0f0f0f0fu, (int)al.z, l1);
                        l0 = dp4a_us(w0.w & 0x0f0f0f0fu, (int)al.w, l0), l1 = dp4a_us(w1.w & 0x0f0f0f0fu, (int)al.w, l1);
                        // high nibbles stay in place (= 16*q as unsigned bytes); their sum is an exact multiple of 16
                        int h0 = dp4a_us(w0.x & 0xf0f0f0f0u, (int)ah.x, 0), h1 = dp4a_us(w1.x & 0xf0f0f0f0u, (int)ah.x, 0);
                        h0 = dp4a_us(w0.y & 0xf0f0f0f0u, (int)ah.y, h0), h1 = dp4a_us(w1.y & 0xf0f0f0f0u, (int)ah.y, h1);
                        h0 = dp4a_us(w0.z & 0xf0f0f0f0u, (int)ah.z, h0), h1 = dp4a_us(w1.z & 0xf0f0f0f0u, (int)ah.z, h1);
                        h0 = dp4a_us(w0.w & 0xf0f0f0f0u, (int)ah.w, h0), h1 = dp4a_us(w1.w & 0xf0f0f0f0u, (int)ah.w, h1);
                        acc0 = fmaf(dw0 * dy, (float)(l0 + (h0 >> 4)), acc0);
                        acc1 = fmaf(dw1 * dy, (float)(l1 + (h1 >> 4)), acc1);
                    }
                }
                __syncwarp();                               // every lane has read the stage: it can be refilled
                if (ig < P.end) {
                    if (lane == 0) issue(st);
                    iadvance();
                }
            }
            const float v0 = warp_sum(acc0), v1 = warp_sum(acc1);
            if (lane == 0) {
                int m = 0;
#pragma unroll
                for (int mm = 1; mm < kPlanMats; ++mm)
                    if (mm < G.nmat && pair >= G.pair_begin[mm]) m = mm;
                const int r0 = 2 * (pair - G.pair_begin[m]);
                float *out = G.out[m];
                const float *res = G.res[m];
                out[r0] = res ? __fadd_rn(v0, __ldcg(res + r0)) : v0;
                if (r0 + 1 < G.rows[m]) out[r0 + 1] = res ? __fadd_rn(v1, __ldcg(res + r0 + 1)) : v1;
            }
        }
    }
    // ---- leave every counter at zero for the next launch
    __syncthreads();
    if (threadIdx.x == 0 && P.end - P.first > 1) {
        const int last = P.end - 1;
        __threadfence();
        const unsigned old = atomicAdd(P.cnt + last, 1u);
        if (old == gridDim.x - 1) {
            if (last - 1 >= P.first) P.cnt[last - 1] = 0;
            P.cnt[last] = 0;
        }
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------------ host side
struct tb200_plan {
    int n_groups, nb_max, device;
    PlanGroup *d_groups;
    unsigned *d_cnt;
    int *d_err;
    size_t smem;
};

extern "C" {

tb200_plan *tb200_plan_create(int n_groups, const tb200_plan_group *groups) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return nullptr;
    if (n_groups < 1 || !groups) { tb200_set_error("plan_create: bad arguments"); return nullptr; }
    PlanGroup *h = (PlanGroup *)calloc((size_t)n_groups, sizeof(PlanGroup));
    if (!h) return nullptr;
    const long GW = (long)c->sm_count * kPW;
    long rot = 0;
    int nb_max = 0;
    for (int g = 0; g < n_groups; ++g) {
        const tb200_plan_group &s = groups[g];
        PlanGroup &G = h[g];
        const char *bad = nullptr;
        if (s.n < 1 || s.n > kPlanMats || !s.x) bad = "1..4 matrices and an input are required";
        else if (s.prologue < TB200_PLAN_QUANT || s.prologue > TB200_PLAN_SWIGLU) bad = "unknown prologue";
        else if (s.prologue == TB200_PLAN_RMSNORM && !s.norm_weight) bad = "RMSNORM needs norm_weight";
        else if (s.prologue == TB200_PLAN_SWIGLU && !s.x2) bad = "SWIGLU needs x2";
        else if (((uintptr_t)s.x | (uintptr_t)s.x2 | (uintptr_t)s.norm_weight) & 15) bad = "inputs must be 16-byte aligned";
        long pairs = 0;
        for (int m = 0; m < s.n && !bad; ++m) {
            const tb200_weights *w = s.w[m];
            if (!w || w->type != TB200_TYPE_Q4_0) bad = "only Q4_0 weights (use the per-matrix API for the other types)";
            else if (w->k != s.w[0]->k || w->k % 256) bad = "matrices of a group share K, a multiple of 256";
            else if (w->rows < 1 || !s.out[m]) bad = "empty matrix or missing output";
            else if (w->device != c->device) bad = "weights live on another device";
            if (bad) break;
            G.pair_begin[m] = (int)pairs;
            pairs += (w->rows + 1) / 2;
            G.rows[m] = (int)w->rows, G.qs[m] = (const uint4 *)w->qs, G.sc[m] = (const __half *)w->sc, G.out[m] = s.out[m], G.res[m] = s.residual[m];
        }
        if (bad) { tb200_set_error("plan_create: group %d: %s", g, bad); free(h); return nullptr; }
        for (int m = s.n; m <= kPlanMats; ++m) G.pair_begin[m] = (int)pairs;
        G.x = s.x, G.x2 = s.x2, G.nw = s.norm_weight, G.eps = s.eps, G.prologue = s.prologue;
        G.nmat = s.n, G.nb = (int)(s.w[0]->k / QK), G.nseg = (G.nb + kSegBlocks - 1) / kSegBlocks;
        G.npairs = (int)pairs, G.rot = (int)(rot % GW);
        rot += pairs;                                       // the pairs left over by one group go to the warps after them in the next
        nb_max = G.nb > nb_max ? G.nb : nb_max;
    }
    tb200_plan *p = new tb200_plan();
    memset(p, 0, sizeof(*p));
    p->n_groups = n_groups, p->nb_max = nb_max, p->device = c->device;
    p->smem = (size_t)kPW * kPS * kStageBytes + (size_t)nb_max * 40;
    bool ok = p->smem <= 220 * 1024;
    if (!ok) tb200_set_error("plan_create: K = %d does not fit the activation staging", nb_max * QK);
    ok = ok && cudaMalloc((void **)&p->d_groups, (size_t)n_groups * sizeof(PlanGroup)) == cudaSuccess;
    ok = ok && cudaMalloc((void **)&p->d_cnt, (size_t)(n_groups + 1) * 4 + 16) == cudaSuccess;
    if (ok) {
        p->d_err = (int *)(p->d_cnt + n_groups + 1);
        ok = cudaMemcpy(p->d_groups, h, (size_t)n_groups * sizeof(PlanGroup), cudaMemcpyHostToDevice) == cudaSuccess &&
             cudaMemset(p->d_cnt, 0, (size_t)(n_groups + 1) * 4 + 16) == cudaSuccess &&
             cudaFuncSetAttribute(plan_q4_0_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p->smem) == cudaSuccess;
        if (!ok) tb200_set_error("plan_create: %s", cudaGetErrorString(cudaGetLastError()));
    } else cudaGetLastError();
    free(h);
    if (!ok) { tb200_plan_free(p); return nullptr; }
    return p;
}

int tb200_plan_run(tb200_plan *p, int first, int count) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    if (!p || first < 0 || count < 0 || first + count > p->n_groups || p->device != c->device) { tb200_set_error("plan_run: bad arguments"); return TB200_ERR_INVALID; }
    if (count == 0) return TB200_OK;
    PlanParams P;
    P.groups = p->d_groups, P.first = first, P.end = first + count, P.cnt = p->d_cnt, P.err = p->d_err, P.nb_max = p->nb_max;
    // one CTA per SM, all co-resident (the groups synchronise through arrival counters); plain stream order, no PDL: the
    // kernel owns the whole GPU for the step
    plan_q4_0_kernel<<<dim3((unsigned)c->sm_count), dim3(kPW * 32), p->smem, c->stream>>>(P);
    tb200_count_launch();
    TB_CUDA(cudaGetLastError());
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
    TB_CUDA(cudaMemset(p->d_cnt, 0, (size_t)(p->n_groups + 1) * 4 + 16));
    return TB200_OK;
}

void tb200_plan_free(tb200_plan *p) {
    if (!p) return;
    if (p->d_groups) cudaFree(p->d_groups);
    if (p->d_cnt) cudaFree(p->d_cnt);
    delete p;
}

}  // extern "C"
