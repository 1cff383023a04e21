// MoE matmul behind GGML_OP_MUL_MAT_ID: expert grouping + row gather.
//
// Reference semantics (llamafile/tinyblas_cpu_mixmul.inc:271-348): quantize every (token, task) row of
// `thought` once; for each expert build the list of (token, thinker) pairs routed to it
// (build_row_pointers :297-320); run the per-expert matmul with B and C addressed through those lists.
// Here the lists are built on the device from `plan` (no host sync, unlike ggml-cuda.cu:2037-2038) as
// fixed-size work items {expert row offset, <=G gathered activation rows, their output rows}; one grouped
// launch serves all experts: decode kernels for small routing counts, tcgen05 GEMM for prefill.
// Plan entries outside [0, experts) match no expert and leave their result row untouched, as in the
// reference's loop.
#include <stdlib.h>
#include <string.h>

#include "tb200_common.cuh"

int tb200_run_matmul(tb200_ctx *c, const tb200_weights *w, long row0, long nrows, const int8_t *aq, const int2 *ads, const void *af,
                     int af_is_f32, long a_stride, long tokens, float *out, long ldo);

int tb200_launch_gemm_grouped(tb200_ctx *c, const tb200_weights *w, int experts, long rows, const int8_t *aq, const int2 *ads, const void *af,
                              const int32_t *plan, long pairs, int thinkers, int tasks, float *result);
bool tb200_gemm_available();

namespace {

// One CTA; warp e (looping if experts > warps) scans all pairs in order, so every expert's list is
// sorted by (token, thinker) exactly like the reference's. Two passes: count, then place.
__global__ void __launch_bounds__(1024) build_items_kernel(const int32_t *__restrict__ plan, long pairs, int thinkers, int tasks,
                                                           int experts, long rows, int group, tb200_gemv_item *__restrict__ items,
                                                           long max_items, int *__restrict__ expert_counts, int *__restrict__ n_live) {
    extern __shared__ int s_base[];   // [experts + 1] first item of each expert
    pdl_launch_dependents();
    pdl_wait();                       // the item table may still be read by the previous call's grouped kernel
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    for (long i = threadIdx.x; i < max_items; i += blockDim.x) items[i].ntok = 0, items[i].row0 = 0;
    for (int e = warp; e < experts; e += nwarps) {
        int count = 0;
        for (long p0 = 0; p0 < pairs; p0 += 32) {
            const long p = p0 + lane;
            const bool hit = p < pairs && plan[p] == e;
            count += __popc(__ballot_sync(0xffffffffu, hit));
        }
        if (lane == 0) { s_base[e + 1] = (count + group - 1) / group; if (expert_counts) expert_counts[e] = count; }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        s_base[0] = 0;
        for (int e = 0; e < experts; ++e) s_base[e + 1] += s_base[e];
        if (n_live) *n_live = s_base[experts];   // live items lead the table (the grouped resident kernel spreads its CTAs over them)
    }
    __syncthreads();
    for (int e = warp; e < experts; e += nwarps) {
        int count = 0;
        for (long p0 = 0; p0 < pairs; p0 += 32) {
            const long p = p0 + lane;
            const bool hit = p < pairs && plan[p] == e;
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            if (hit) {
                const int j = count + __popc(m & ((1u << lane) - 1));
                tb200_gemv_item &it = items[s_base[e] + j / group];
                const long token = p / thinkers;
                const int thinker = (int)(p % thinkers);
                it.a_row[j % group] = (int)(token * tasks + thinker % tasks);   // thinker % tasks: tinyblas_cpu_mixmul.inc:311-317
                it.out_row[j % group] = (int)p;
            }
            count += __popc(m);
        }
        if (lane == 0)
            for (int j0 = 0; j0 < count; j0 += group) {
                tb200_gemv_item &it = items[s_base[e] + j0 / group];
                it.row0 = (long)e * rows;
                it.ntok = min(group, count - j0);
            }
    }
}

size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

template <int G>
int grouped(tb200_ctx *c, const tb200_weights *w, long rows, const int8_t *aq, const int2 *ads, const void *af, long k,
            const tb200_gemv_item *items, long n_items, const int *n_live, float *result) {
    tb200_gemv_args a;
    memset(&a, 0, sizeof(a));
    a.nmat = 1;
    a.w = w, a.row0 = 0, a.nrows = rows, a.aq = aq, a.ads = ads, a.af = af, a.af_is_f32 = 0, a.a_stride = k;
    a.tokens = G, a.out = result, a.ldo = rows, a.items = items, a.n_items = n_items, a.n_live = n_live;
    return tb200_launch_gemv(c, a);
}

}  // namespace

extern "C" int tb200_mixmul_dev(const tb200_weights *w, int experts, const float *thought, int tasks, const int32_t *plan,
                                int thinkers, long tokens, float *result) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    if (!w || experts <= 0 || w->rows % experts || tasks <= 0 || tasks > thinkers || thinkers > experts || tokens < 0) {
        tb200_set_error("mixmul: bad arguments");
        return TB200_ERR_INVALID;
    }
    // weight types MixMul::mixmuler serves (tinyblas_cpu_mixmul.inc:162-269), minus F32/BF16 which this module does not hold
    if (w->type != TB200_TYPE_Q4_0 && w->type != TB200_TYPE_Q8_0 && w->type != TB200_TYPE_F16) return TB200_ERR_UNSUPPORTED;
    const long rows = w->rows / experts, k = w->k, pairs = tokens * thinkers, arows = tokens * tasks;
    if (k % (w->type == TB200_TYPE_F16 ? 16 : 32)) return TB200_ERR_UNSUPPORTED;   // mixmat: cols % KN (:273)
    if (pairs == 0 || rows == 0) return TB200_OK;
    // tokens per work item. Q4_0 has the grouped resident kernel, whose cost follows the item's real token count, so it takes
    // the widest group whose staged activations still leave two CTAs per SM (fewer items = fewer passes over an expert)
    int group = pairs <= experts ? 1 : (pairs <= 4L * experts ? 4 : 8);
    if (w->type == TB200_TYPE_Q4_0 && group > 1) group = (size_t)(k / QK) * (8 * 40 + 16) <= 110 * 1024 ? 8 : 4;   // two CTAs per SM at least
    const long max_items = group == 1 ? pairs : (pairs + group - 1) / group + (pairs < experts ? pairs : experts);   // sum_e ceil(c_e / group)

    // scratch: planar activations | items
    size_t q_bytes = 0, ds_bytes = 0, h_bytes = 0;
    if (w->type == TB200_TYPE_F16) h_bytes = round_up((size_t)arows * k * 2, 256);
    else q_bytes = round_up((size_t)arows * k, 256), ds_bytes = round_up((size_t)arows * (k / QK) * 8, 256);
    const size_t items_bytes = round_up((size_t)max_items * sizeof(tb200_gemv_item), 256) + 256;   // + live-item count
    uint8_t *base = (uint8_t *)tb200_scratch(c, q_bytes + ds_bytes + h_bytes + items_bytes);
    if (!base) return TB200_ERR_CUDA;
    int8_t *aq = (int8_t *)base;
    int2 *ads = (int2 *)(base + q_bytes);
    __half *ah = (__half *)(base + q_bytes + ds_bytes);
    tb200_gemv_item *items = (tb200_gemv_item *)(base + q_bytes + ds_bytes + h_bytes);
    int *n_live = (int *)(base + q_bytes + ds_bytes + h_bytes + items_bytes - 256);

    int rc;
    if (w->type == TB200_TYPE_F16) rc = tb200_launch_f32_to_f16(c, thought, k, arows, k, ah);          // quantize_thought :322-348
    else rc = tb200_launch_quantize_q8_0(c, thought, k, arows, k, nullptr, aq, ads);
    if (rc) return rc;
    // prefill-sized routing: one grouped tcgen05 launch over all experts
    static const long gemm_from = [] { const char *v = getenv("TB200_MIXMUL_GEMM_PAIRS"); return v && *v ? atol(v) : 16L; }();   // routed pairs per expert
    if (pairs >= gemm_from * experts && k % 256 == 0 && tb200_gemm_available())
        return tb200_launch_gemm_grouped(c, w, experts, rows, w->type == TB200_TYPE_F16 ? nullptr : aq, w->type == TB200_TYPE_F16 ? nullptr : ads,
                                         w->type == TB200_TYPE_F16 ? (const void *)ah : nullptr, plan, pairs, thinkers, tasks, result);
    const int threads = experts >= 32 ? 1024 : experts * 32;
    TB_CUDA(tb200_launch(build_items_kernel, dim3(1), dim3(threads), (experts + 1) * sizeof(int), c->stream, plan, pairs, thinkers, tasks, experts, rows,
                         group, items, max_items, (int *)nullptr, n_live));
    const void *af = w->type == TB200_TYPE_F16 ? (const void *)ah : nullptr;
    if (w->type == TB200_TYPE_F16) aq = nullptr, ads = nullptr;
    switch (group) {
    case 1: return grouped<1>(c, w, rows, aq, ads, af, k, items, max_items, nullptr, result);
    case 4: return grouped<4>(c, w, rows, aq, ads, af, k, items, max_items, n_live, result);
    default: return grouped<8>(c, w, rows, aq, ads, af, k, items, max_items, n_live, result);
    }
}
