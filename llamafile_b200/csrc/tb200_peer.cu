// Multi-GPU exchange for the row-sharded decode path, without a collective library call on the data
// path: the GEMV epilogue of every rank stores its output rows straight into every rank's copy of a symmetric output
// buffer over NVLink (tb200_gemv.cu store_rows), so "all-gather" is just those stores plus the flag barrier below.
// The reference moves these slices with cudaMemcpyPeerAsync + events per device (ggml-cuda.cu:1600-1676).
// One process per GPU; buffers are shared with CUDA IPC handles that the host exchanges (torch.distributed in bench/tests).
#include <string.h>

#include "tb200_common.cuh"

namespace {

struct BarrierArgs {
    uint32_t *flags[TB200_MAX_PEERS];
    uint32_t *epoch;
    int *timeout_flag;
    int world, rank;
};

// Launched with programmatic serialization like every kernel of the module, so it is already resident when the preceding
// GEMV drains; pdl_wait() returns once that GEMV has completed, i.e. after all of this rank's peer stores have been
// issued and flushed. Thread p then publishes this rank's epoch to rank p and waits for rank p's epoch to arrive here.
// The wait is bounded (~2 s) so a missing peer cannot wedge the GPU.
__global__ void peer_barrier_kernel(const BarrierArgs a) {
    pdl_launch_dependents();
    pdl_wait();
    __shared__ uint32_t e;
    if (threadIdx.x == 0) e = ++(*a.epoch);
    __syncthreads();
    const int p = threadIdx.x;
    if (p >= a.world) return;
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a.flags[p] + a.rank), "r"(e) : "memory");
    const uint32_t *mine = a.flags[a.rank] + p;
    const long long t0 = clock64();
    uint32_t v;
    do {
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
        if ((int32_t)(v - e) >= 0) break;
        if (clock64() - t0 > 4000000000LL) { *a.timeout_flag = 1; break; }
    } while (true);
}

}  // namespace

extern "C" {

int tb200_ipc_get_handle(const void *dev_ptr, void *handle64) {
    if (!tb200_get_ctx()) return TB200_ERR_NO_DEVICE;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaIpcMemHandle_t h;
    TB_CUDA(cudaIpcGetMemHandle(&h, const_cast<void *>(dev_ptr)));
    memcpy(handle64, &h, 64);
    return TB200_OK;
}

void *tb200_ipc_open_handle(const void *handle64) {
    if (!tb200_get_ctx()) return nullptr;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    void *p = nullptr;
    if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
        tb200_set_error("ipc_open_handle: %s", cudaGetErrorString(cudaGetLastError()));
        return nullptr;
    }
    return p;
}

int tb200_ipc_close_handle(void *p) {
    if (p) TB_CUDA(cudaIpcCloseMemHandle(p));
    return TB200_OK;
}

tb200_peers *tb200_peers_create(int world, int rank, void *const *out_bases, void *const *flag_bases) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return nullptr;
    if (world < 1 || world > TB200_MAX_PEERS || rank < 0 || rank >= world || !out_bases || !flag_bases) { tb200_set_error("peers_create: bad arguments"); return nullptr; }
    tb200_peers *g = new tb200_peers();
    memset(g, 0, sizeof(*g));
    g->world = world, g->rank = rank;
    for (int p = 0; p < world; ++p) g->out_base[p] = out_bases[p], g->flags[p] = (uint32_t *)flag_bases[p];
    if (cudaMalloc(&g->epoch, 16) != cudaSuccess || cudaMemsetAsync(g->epoch, 0, 16, c->stream) != cudaSuccess) { cudaGetLastError(); delete g; return nullptr; }
    g->timeout_flag = (int *)(g->epoch + 1);
    g->plan_seq = g->epoch + 2;
    return g;
}

void tb200_peers_free(tb200_peers *g) {
    if (!g) return;
    if (g->epoch) cudaFree(g->epoch);
    delete g;
}

int tb200_peers_barrier(tb200_peers *g) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return TB200_ERR_NO_DEVICE;
    if (!g) return TB200_ERR_INVALID;
    BarrierArgs a;
    for (int p = 0; p < TB200_MAX_PEERS; ++p) a.flags[p] = g->flags[p];
    a.epoch = g->epoch, a.timeout_flag = g->timeout_flag, a.world = g->world, a.rank = g->rank;
    TB_CUDA(tb200_launch(peer_barrier_kernel, dim3(1), dim3(32), 0, c->stream, a));
    return TB200_OK;
}

int tb200_peers_timed_out(tb200_peers *g) {
    int v = 0;
    if (g && cudaMemcpy(&v, g->timeout_flag, 4, cudaMemcpyDeviceToHost) != cudaSuccess) { cudaGetLastError(); return -1; }
    return v;
}

}  // extern "C"
