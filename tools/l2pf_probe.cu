// Micro-benchmark behind the decode plan's prefetch design: 148 x 15 warps stream a large buffer through per-warp shared-memory
// rings filled with 1-D bulk copies (the plan kernel's access pattern), optionally with a cp.async.bulk.prefetch.L2 issued P
// items ahead of every copy. Prints GB/s per variant; run under ncu for dram__bytes_read to see whether prefetched lines
// survive until the copy asks for them.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_variants/l2pf_probe tools/l2pf_probe.cu && tools/_variants/l2pf_probe
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

constexpr int kWarps = 15, kStages = 2, kItem = 4608;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(kWarps * 32, 1) stream_kernel(const uint8_t *buf, long items_per_warp, int P, int hint, int stall_every, int stall_cycles,
                                                                unsigned long long *sink, unsigned long long *issue_cyc, int active_warps) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t bars[kWarps][kStages];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long GW = (long)gridDim.x * kWarps, gw = (long)blockIdx.x * kWarps + warp;
    if (threadIdx.x == 0) {
        for (int i = 0; i < kWarps * kStages; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[0][0] + i)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (warp >= active_warps) return;
    long long t_exp = 0, t_cp = 0;
    uint64_t policy;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
    uint8_t *ring = smem + (size_t)warp * kStages * kItem;
    auto src = [&](long i) { return buf + ((size_t)i * GW + gw) * kItem; };   // item i of this warp: consecutive warps read consecutive items
    auto issue = [&](long i) {
        if (i >= items_per_warp) return;
        const int st = (int)(i % kStages);
        if (lane == 0) {
            const long long a0 = clock64();
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bars[warp][st])), "r"(kItem) : "memory");
            const long long a1 = clock64();
            t_exp += a1 - a0;
            if (hint)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                             ::"r"(smem_u32(ring + st * kItem)), "l"(src(i)), "r"(kItem), "r"(smem_u32(&bars[warp][st])), "l"(policy) : "memory");
            else
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(smem_u32(ring + st * kItem)), "l"(src(i)), "r"(kItem), "r"(smem_u32(&bars[warp][st])) : "memory");
            t_cp += clock64() - a1;
            if (P > 0 && i + P < items_per_warp) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src(i + P)), "r"(kItem) : "memory");
        }
    };
    if (lane == 0 && P > 0)
        for (long i = kStages; i < kStages + P && i < items_per_warp; ++i) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src(i)), "r"(kItem) : "memory");
    for (int s = 0; s < kStages; ++s) issue(s);
    unsigned acc = 0;
    long long t_issue = 0, t_wait = 0, t_use = 0;
    for (long i = 0; i < items_per_warp; ++i) {
        const long long c0 = clock64();
        const int st = (int)(i % kStages);
        const uint32_t parity = (uint32_t)((i / kStages) & 1);
        uint32_t ok = 0;
        while (!ok)
            asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(ok) : "r"(smem_u32(&bars[warp][st])), "r"(parity) : "memory");
        const long long c1 = clock64();
        const uint4 *p = reinterpret_cast<const uint4 *>(ring + st * kItem);
        for (int j = lane; j < kItem / 16; j += 32) { const uint4 v = p[j]; acc += v.x ^ v.y ^ v.z ^ v.w; }
        __syncwarp();
        const long long c2 = clock64();
        issue(i + kStages);
        const long long c3 = clock64();
        t_wait += c1 - c0, t_use += c2 - c1, t_issue += c3 - c2;
        if (stall_every > 0 && (i + 1) % stall_every == 0) {     // emulate the plan's group boundary: nobody consumes for a while
            const long long t0 = clock64();
            while (clock64() - t0 < stall_cycles) {}
        }
    }
    if (acc == 0x12345678u) *sink = acc;
    if (threadIdx.x == 0 && blockIdx.x == 7) issue_cyc[0] = t_wait / items_per_warp, issue_cyc[1] = t_use / items_per_warp, issue_cyc[2] = t_issue / items_per_warp, issue_cyc[3] = t_exp / items_per_warp, issue_cyc[4] = t_cp / items_per_warp;
}

int main(int argc, char **argv) {
    const size_t bytes = (size_t)(argc > 1 ? atof(argv[1]) : 4.0) * (1ull << 30);
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    uint8_t *buf;
    unsigned long long *sink, *cyc; cudaMallocManaged(&cyc, 64);
    cudaMalloc(&buf, bytes);
    cudaMalloc(&sink, 8);
    cudaMemset(buf, 1, bytes);
    const long GW = (long)sms * kWarps, items_per_warp = (long)(bytes / kItem / GW);
    const size_t smem = (size_t)kWarps * kStages * kItem;
    cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0), cudaEventCreate(&e1);
    const int only = argc > 2 ? atoi(argv[2]) : -1;      // run a single variant (for ncu)
    struct V { int P, hint, stall_every, stall_cycles, aw; } vs[] = {
        // L2 prefetch P items ahead of every copy, with and without the evict-first hint on the copies
        {0, 1, 0, 0, 15}, {0, 0, 0, 0, 15}, {1, 1, 0, 0, 15}, {2, 1, 0, 0, 15}, {4, 1, 0, 0, 15}, {8, 1, 0, 0, 15}, {4, 0, 0, 0, 15},
        // the same with every warp pausing ~3 us after every 2 / 6 items (a group boundary of the decode plan)
        {0, 1, 2, 6000, 15}, {2, 1, 2, 6000, 15}, {4, 1, 2, 6000, 15}, {8, 1, 2, 6000, 15}, {0, 1, 6, 6000, 15}, {4, 1, 6, 6000, 15}, {8, 1, 6, 6000, 15}, {16, 1, 6, 6000, 15},
        // cost of issuing a copy (expect_tx + cp.async.bulk) at different load levels: 8, 4, 2, 1 active warps per SM
        {0, 1, 0, 0, 8}, {0, 1, 0, 0, 4}, {0, 1, 0, 0, 2}, {0, 1, 0, 0, 1},
    };
    int idx = 0;
    for (const V &v : vs) {
        if (only >= 0 && idx++ != only) continue;
        float best = 1e30f;
        for (int rep = 0; rep < (only >= 0 ? 1 : 3); ++rep) {
            cudaEventRecord(e0);
            stream_kernel<<<sms, kWarps * 32, smem>>>(buf, items_per_warp, v.P, v.hint, v.stall_every, v.stall_cycles, sink, cyc, v.aw);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            best = ms < best ? ms : best;
        }
        const double moved = (double)items_per_warp * GW * kItem * v.aw / kWarps;
        // with stalls: the ideal is max(stream time, stall time); report both
        const double stall_ms = v.stall_every ? (double)(items_per_warp / v.stall_every) * v.stall_cycles / 1.965e6 : 0.0;
        printf("P=%2d hint=%d stall %d items/%d cyc: %8.3f ms  %7.1f GB/s  (stalls alone %.3f ms, stream alone %.3f ms at 6500 GB/s) %s\n", v.P, v.hint, v.stall_every,
               v.stall_cycles, best, moved / best / 1e6, stall_ms, moved / 6.5e9, cudaGetErrorString(cudaGetLastError()));
        printf("      active warps %d; per item (warp 0 of CTA 7): wait %llu  use %llu  issue %llu (expect_tx %llu, bulk copy %llu) cycles\n", v.aw, cyc[0], cyc[1], cyc[2], cyc[3], cyc[4]);
    }
    return 0;
}
