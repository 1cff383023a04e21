// The reference's own entry points on HOST pointers (include/tinyblas_b200.h part 1):
//   llamafile_sgemm, llamafile_mixmul, llamafile_mixmul_needs, llamafile_mixmul_iqk  (llamafile/sgemm.h:23-28, 76-77)
//   quantize_row_q8_0 / quantize_row_q8_K / ggml_fp32_to_fp16_row as tb200_* (from_float of the vec_dot types)
// Same argument meaning and true/false behaviour as the CPU implementation; the work itself runs on the
// GPU: weights are uploaded once and cached by host address, activations go H2D, results come D2H, and the
// call returns when the result is in the caller's buffer. Nothing here computes on the CPU.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <pthread.h>

#include <atomic>
#include <map>
#include <mutex>
#include <tuple>
#include <vector>

#include "tb200_common.cuh"

int tb200_run_matmul(tb200_ctx *c, const tb200_weights *w, long row0, long nrows, const int8_t *aq, const int2 *ads, const void *af,
                     int af_is_f32, long a_stride, long tokens, float *out, long ldo);
bool tb200_serviced(int Atype, int Btype, int Ctype, long n);
size_t tb200_gguf_row_bytes(int type, long k);

namespace {

std::mutex g_cache_mu;
std::map<std::tuple<const void *, long, long, size_t, int>, tb200_weights *> g_cache;
// Host ranges the caller has declared immutable for the lifetime of the registration (the mmap'd GGUF file,
// tb200_register_weights): only matrices lying inside one are uploaded once and kept. Any other A — the F16 or quantized
// KV cache that ggml also sends through llamafile_sgemm (llama.cpp/src/llama.cpp:8387/8430), scratch tensors — is
// uploaded on every call and dropped after it, because its bytes may change between calls at the same address.
std::vector<std::pair<const char *, size_t>> g_ranges;

// Sticky failure state. The reference's true/false answer depends only on the arguments, so every ggml worker gets the
// same one; here threads ith != 0 return true before thread 0 has touched the GPU. A runtime failure on thread 0
// (allocation, launch, CUDA error) therefore cannot be turned into "false" for everybody after the fact: it is fatal for
// the call in flight, like the reference's asserts (tinyblas_cpu_sgemm.inc:277-284) — logged and abort()ed unless
// TB200_ABORT_ON_ERROR=0 — and the module answers false on EVERY thread from then on.
std::atomic<bool> g_disabled{false};
bool runtime_failure(const char *where, int nth) {
    g_disabled.store(true, std::memory_order_release);
    fprintf(stderr, "tinyblas_b200: runtime failure in %s: %s — module disabled, later calls return false on every thread\n", where, tb200_last_error());
    static const bool hard = [] { const char *v = getenv("TB200_ABORT_ON_ERROR"); return !(v && *v == '0'); }();
    if (hard && nth > 1) abort();   // with one caller thread nobody has answered true yet: "false" is still a consistent answer
    return false;
}

// One host-pointer call at a time per device: the lane (stream, staging pages, graphs), the io buffer and the context's
// stream swap below are per-device state. The lock is only taken by the thread that does the work (ith == 0); ggml
// workers may be pthread_cancel()led, so cancellation is held off while the lock is owned.
std::mutex g_dev_mu[16];
struct DeviceLock {
    std::mutex &mu; int old_state = 0;
    explicit DeviceLock(int device) : mu(g_dev_mu[device & 15]) { pthread_setcancelstate(PTHREAD_CANCEL_DISABLE, &old_state); mu.lock(); }
    ~DeviceLock() { mu.unlock(); pthread_setcancelstate(old_state, nullptr); }
};

// device staging for host-pointer calls, separate from the kernels' scratch arena
struct Io { void *p = nullptr; size_t cap = 0; };
Io g_io[16];

void *io_buffer(tb200_ctx *c, size_t bytes) {
    Io &io = g_io[c->device];
    if (bytes <= io.cap) return io.p;
    cudaStreamSynchronize(c->stream);
    if (io.p) cudaFree(io.p);
    io.p = nullptr, io.cap = 0;
    const size_t want = (bytes + bytes / 2 + (1 << 20) - 1) >> 20 << 20;
    if (cudaMalloc(&io.p, want) != cudaSuccess) { cudaGetLastError(); tb200_set_error("io: cudaMalloc(%zu) failed", want); return nullptr; }
    io.cap = want;
    return io.p;
}

size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }

// ---- decode lane: host-mapped staging, completion word, cached launch graphs ---------------------------------------
// A decode-size call (n <= 8) moves a few KB each way, so its cost is latency, not bandwidth. Measured on the B200
// box (tools/host_latency.py): every cudaLaunchKernelEx costs ~3.3 us of host time and a kernel starts several us
// after its launch call, so "H2D copy, two kernels, D2H copy, cudaStreamSynchronize" is ~30 us around ~4 us of GPU
// work. The lane cuts that down:
//   * activation blocks are read by the unpack kernel straight from host-MAPPED pinned memory (the caller's own
//     buffer when it is already pinned, else one memcpy into the lane's staging page): no copy-engine transfer;
//   * a small C is stored by the decode kernel straight into mapped host memory; a large one (scattered 4-byte
//     PCIe stores cost ~0.6 us per 1000 rows) goes to HBM and leaves through one copy-engine transfer;
//   * completion is a sequence number published by a one-thread kernel (st.release.sys after griddepcontrol.wait,
//     i.e. after every store of its predecessors is performed); the caller spins on that word instead of
//     cudaStreamSynchronize();
//   * the launches of one (weights, n, buffers) combination are captured once into a CUDA graph and replayed with a
//     single cudaGraphLaunch on the lane's own stream (the entry points are synchronous, so nothing else orders on it).
struct LaneKey {
    const tb200_weights *w; long n; int btype; const void *dB; void *dC; long ldb, ldo; void *h_dst;
    bool operator<(const LaneKey &o) const {
        return std::tie(w, n, btype, dB, dC, ldb, ldo, h_dst) < std::tie(o.w, o.n, o.btype, o.dB, o.dC, o.ldb, o.ldo, o.h_dst);
    }
};
struct HostLane {
    uint8_t *in = nullptr;  size_t in_cap = 0;     // mapped pinned, host address == device address (UVA)
    uint8_t *out = nullptr; size_t out_cap = 0;
    uint8_t *dact = nullptr; size_t dact_cap = 0;  // device: activations in the kernels' planar form
    uint8_t *dout = nullptr; size_t dout_cap = 0;  // device: large C before its D2H transfer
    uint32_t *flag = nullptr;                      // mapped pinned completion word
    uint32_t *dcount = nullptr;                    // device: publish counter (graph replays advance it); dcount[1] = CTA arrivals of the copy-out
    uint32_t seq = 0;                              // host mirror of dcount
    cudaStream_t stream = nullptr;
    std::map<LaneKey, cudaGraphExec_t> graphs;     // exec == nullptr: seen once, not captured yet
};
HostLane g_lane[16];
// C goes through HBM and is written out by copy_publish_kernel in 512-byte warp stores. TB200_LANE_DIRECT_BYTES=n lets
// the decode kernel store a C of up to n bytes straight into host memory instead (scattered 4-byte PCIe stores, ~0.6 us per
// 1000 rows: N=4096 17.9 us direct vs 15.5 us copied, N=1024 equal; whole-step e2e 924 GB/s with n=4096 vs 978 GB/s with 0).
// Beyond kLaneKernelCopyBytes the copy engine takes over.
size_t lane_direct_bytes() { static const size_t v = [] { const char *e = getenv("TB200_LANE_DIRECT_BYTES"); return e && *e ? (size_t)atol(e) : (size_t)0; }(); return v; }
constexpr size_t kLaneKernelCopyBytes = 4u << 20;

__global__ void publish_kernel(uint32_t *flag, uint32_t *count) {
    pdl_launch_dependents();
    pdl_wait();                                    // predecessor grid complete, its stores performed
    const uint32_t v = *count + 1;
    *count = v;
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(v) : "memory");
}
// Large C: HBM -> mapped pinned host memory with coalesced stores, then the completion word, in ONE kernel (a
// cudaMemcpyAsync node between the decode kernel and publish_kernel cost ~15 us of copy-engine latency per call: N=14336
// took 37 us against 18 us for N=4096). Small results use one CTA (its thread 0 publishes with a release store after the CTA barrier); larger ones
// use several, thread 0 of each fences the CTA's stores system-wide and counts the CTA in, and the last to arrive publishes.
__global__ void __launch_bounds__(1024) copy_publish_kernel(const float *__restrict__ src, float *__restrict__ dst, long n, uint32_t *flag, uint32_t *count) {
    pdl_launch_dependents();
    pdl_wait();
    const long tid = blockIdx.x * (long)blockDim.x + threadIdx.x, nthr = (long)gridDim.x * blockDim.x;
    if (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0) {
        const long n4 = n >> 2;
        const uint4 *s4 = reinterpret_cast<const uint4 *>(src);
        uint4 *d4 = reinterpret_cast<uint4 *>(dst);
        for (long i = tid; i < n4; i += nthr) d4[i] = s4[i];
        for (long i = (n4 << 2) + tid; i < n; i += nthr) dst[i] = src[i];
    } else {
        for (long i = tid; i < n; i += nthr) dst[i] = src[i];
    }
    __syncthreads();                 // the CTA's stores are ordered before thread 0's release (cumulative), so ONE system-scope
    if (threadIdx.x == 0) {          // fence per CTA is enough: a fence in every thread cost ~9 us per call
        if (gridDim.x > 1) {
            __threadfence_system();
            if (atomicAdd(count + 1, 1u) != gridDim.x - 1) return;
            count[1] = 0;
        }
        const uint32_t v = *count + 1;
        *count = v;
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(v) : "memory");
    }
}
// fp16 / f32 activations of the F16 path: every CTA re-reads them, so they must come from HBM, not over PCIe
__global__ void stage_copy_kernel(const uint4 *__restrict__ src, uint4 *__restrict__ dst, long n16, const uint8_t *src_tail, uint8_t *dst_tail, int tail) {
    pdl_launch_dependents();
    pdl_wait();
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n16; i += (long)gridDim.x * blockDim.x) dst[i] = src[i];
    if (blockIdx.x == 0 && (int)threadIdx.x < tail) dst_tail[threadIdx.x] = src_tail[threadIdx.x];
}

void lane_drop_graphs(HostLane &L) {
    for (auto &kv : L.graphs)
        if (kv.second) cudaGraphExecDestroy(kv.second);
    L.graphs.clear();
}

bool lane_reserve_host(HostLane &L, uint8_t **p, size_t *cap, size_t bytes) {
    if (bytes <= *cap) return true;
    lane_drop_graphs(L);   // captured graphs hold the old address
    if (*p) { cudaDeviceSynchronize(); cudaFreeHost(*p); *p = nullptr; *cap = 0; }
    const size_t want = (bytes + bytes / 2 + 65535) & ~(size_t)65535;
    if (cudaHostAlloc((void **)p, want, cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError(); tb200_set_error("lane: cudaHostAlloc(%zu) failed", want); *p = nullptr; return false;
    }
    *cap = want;
    return true;
}
bool lane_reserve_dev(HostLane &L, uint8_t **p, size_t *cap, size_t bytes) {
    if (bytes <= *cap) return true;
    lane_drop_graphs(L);
    if (*p) { cudaDeviceSynchronize(); cudaFree(*p); *p = nullptr; *cap = 0; }
    const size_t want = (bytes + bytes / 2 + 65535) & ~(size_t)65535;
    if (cudaMalloc((void **)p, want) != cudaSuccess) { cudaGetLastError(); tb200_set_error("lane: cudaMalloc(%zu) failed", want); *p = nullptr; return false; }
    *cap = want;
    return true;
}
bool lane_init(HostLane &L) {
    if (L.flag) return true;
    if (cudaStreamCreateWithFlags(&L.stream, cudaStreamNonBlocking) != cudaSuccess) { cudaGetLastError(); return false; }
    if (cudaMalloc((void **)&L.dcount, 64) != cudaSuccess || cudaMemset(L.dcount, 0, 64) != cudaSuccess) { cudaGetLastError(); return false; }
    if (cudaHostAlloc((void **)&L.flag, 64, cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return false; }
    *L.flag = 0;
    return true;
}

// device-usable alias of a host pointer that is already pinned (cudaHostAlloc / cudaHostRegister), else nullptr
void *pinned_alias(const void *h) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, h) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return at.type == cudaMemoryTypeHost ? at.devicePointer : nullptr;
}

inline void cpu_relax() {
#if defined(__x86_64__) || defined(__i386__)
    __builtin_ia32_pause();
#elif defined(__aarch64__)
    asm volatile("yield" ::: "memory");
#endif
}

// spin on the completion word; every ~4k spins look at the stream so a failed launch cannot hang the caller
bool lane_wait(HostLane &L, uint32_t seq) {
    volatile uint32_t *f = L.flag;
    for (unsigned long spins = 1;; ++spins) {
        if (*f == seq) { __atomic_thread_fence(__ATOMIC_ACQUIRE); return true; }
        if ((spins & 0xfff) == 0) {
            const cudaError_t q = cudaStreamQuery(L.stream);
            if (q == cudaSuccess) { __atomic_thread_fence(__ATOMIC_ACQUIRE); return *f == seq; }
            if (q != cudaErrorNotReady) { cudaGetLastError(); tb200_set_error("lane: stream error %s", cudaGetErrorString(q)); return false; }
        }
        cpu_relax();
    }
}

// TB200_LANE_TRACE=1: cumulative wall time of the lane's phases, printed every 1024 calls (debug aid)
struct LaneTrace { double prep = 0, launch = 0, wait = 0, out = 0; long calls = 0, replays = 0; };
LaneTrace g_trace;
bool env_flag(const char *name, int dflt) { const char *v = getenv(name); return v && *v ? *v != '0' : dflt != 0; }
bool trace_enabled() { static const bool on = env_flag("TB200_LANE_TRACE", 0); return on; }
bool lane_enabled() { static const bool on = env_flag("TB200_HOST_LANE", 1); return on; }
bool lane_graphs_enabled() { static const bool on = env_flag("TB200_LANE_GRAPH", 1); return on; }
double now_us() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3; }

// the launches of one decode call, on c->stream (which the caller has pointed at the lane's stream)
int lane_issue(tb200_ctx *c, HostLane &L, const LaneKey &k, size_t b_bytes) {
    const tb200_weights *w = k.w;
    int rc;
    if (w->type == TB200_TYPE_F16) {
        const long n16 = (long)(b_bytes / 16);
        const unsigned grid = (unsigned)((n16 + 255) / 256 > 0 ? (n16 + 255) / 256 : 1);
        TB_CUDA(tb200_launch(stage_copy_kernel, dim3(grid < 1184u ? grid : 1184u), dim3(256), 0, c->stream, (const uint4 *)k.dB, (uint4 *)L.dact, n16,
                             (const uint8_t *)k.dB + n16 * 16, L.dact + n16 * 16, (int)(b_bytes & 15)));
        rc = tb200_run_matmul(c, w, 0, w->rows, nullptr, nullptr, L.dact, k.btype == TB200_TYPE_F32, k.ldb, k.n, (float *)k.dC, k.ldo);
    } else {
        int8_t *q = (int8_t *)L.dact;
        int2 *ds = (int2 *)(L.dact + up256((size_t)k.n * w->k));
        rc = tb200_is_kquant(w->type) ? tb200_launch_unpack_q8_K(c, k.dB, k.ldb, k.n, w->k, q, ds)
             : k.btype == TB200_TYPE_Q8_1 ? tb200_launch_unpack_q8_1(c, k.dB, k.ldb, k.n, w->k, q, ds)
                                          : tb200_launch_unpack_q8_0(c, k.dB, k.ldb, k.n, w->k, q, ds);
        if (rc) return rc;
        rc = tb200_run_matmul(c, w, 0, w->rows, q, ds, nullptr, 0, w->k, k.n, (float *)k.dC, k.ldo);
    }
    if (rc) return rc;
    if (k.h_dst) {   // large C: HBM -> pinned host (dense rows)
        const size_t c_bytes = (size_t)k.n * w->rows * 4;
        void *h_dev = c_bytes <= kLaneKernelCopyBytes ? pinned_alias(k.h_dst) : nullptr;
        if (h_dev) {
            const long n = (long)(c_bytes / 4);
            // up to 128 KB: ONE CTA, so one fence and no arrival counting (two dependent system fences cost ~4 us)
            long ctas = c_bytes <= 128 * 1024 ? 1 : (n / 4 + 1023) / 1024;
            ctas = ctas > c->sm_count ? c->sm_count : ctas;
            TB_CUDA(tb200_launch(copy_publish_kernel, dim3((unsigned)ctas), dim3(1024), 0, c->stream, (const float *)k.dC, (float *)h_dev, n, L.flag, L.dcount));
            return TB200_OK;
        }
        TB_CUDA(cudaMemcpyAsync(k.h_dst, k.dC, c_bytes, cudaMemcpyDeviceToHost, c->stream));
    }
    TB_CUDA(tb200_launch(publish_kernel, dim3(1), dim3(1), 0, c->stream, L.flag, L.dcount));
    return TB200_OK;
}

// Resident copy of A. *transient is set when A lies outside every registered range: the caller frees the copy after the call.
tb200_weights *cached_weights(const void *A, int type, long rows, long k, size_t row_stride, bool *transient) {
    std::lock_guard<std::mutex> lk(g_cache_mu);
    const size_t rb = tb200_gguf_row_bytes(type, k);
    const size_t span = rows ? (size_t)(rows - 1) * (row_stride ? row_stride : rb) + rb : 0;
    bool inside = false;
    for (const auto &r : g_ranges)
        if ((const char *)A >= r.first && (const char *)A + span <= r.first + r.second) { inside = true; break; }
    *transient = !inside;
    if (!inside) return tb200_weights_upload(A, type, rows, k, row_stride);
    const auto key = std::make_tuple(A, rows, k, row_stride, type);
    auto it = g_cache.find(key);
    if (it != g_cache.end()) return it->second;
    tb200_weights *w = tb200_weights_upload(A, type, rows, k, row_stride);
    if (w) g_cache[key] = w;
    return w;
}
struct WeightsRef {   // frees a transient copy when the call ends
    tb200_weights *w = nullptr; bool transient = false;
    ~WeightsRef() { if (w && transient) tb200_weights_free(w); }
};

int blck(int type) {
    if (tb200_is_kquant(type) || type == TB200_TYPE_Q8_K) return QK_K;
    return (type == TB200_TYPE_Q4_0 || type == TB200_TYPE_Q8_0 || type == TB200_TYPE_Q4_1 || type == TB200_TYPE_Q5_0 || type == TB200_TYPE_Q8_1) ? QK : 1;
}
size_t tsize(int type) { return tb200_gguf_row_bytes(type, blck(type)); }

}  // namespace

extern "C" {

static void drop_cached_locked(const char *lo, const char *hi) {   // g_cache_mu held; nullptr range = everything
    for (HostLane &L : g_lane) {   // captured graphs hold weight pointers
        if (L.stream) cudaStreamSynchronize(L.stream);
        lane_drop_graphs(L);
    }
    for (auto it = g_cache.begin(); it != g_cache.end();) {
        const char *a = (const char *)std::get<0>(it->first);
        if (!lo || (a >= lo && a < hi)) { tb200_weights_free(it->second); it = g_cache.erase(it); }
        else ++it;
    }
}

void tb200_weights_cache_clear(void) {
    std::lock_guard<std::mutex> lk(g_cache_mu);
    drop_cached_locked(nullptr, nullptr);
}

int tb200_register_weights(const void *base, size_t bytes) {
    if (!base || !bytes) { tb200_set_error("register_weights: empty range"); return TB200_ERR_INVALID; }
    std::lock_guard<std::mutex> lk(g_cache_mu);
    g_ranges.emplace_back((const char *)base, bytes);
    return TB200_OK;
}

int tb200_unregister_weights(const void *base) {
    std::lock_guard<std::mutex> lk(g_cache_mu);
    for (size_t i = 0; i < g_ranges.size(); ++i)
        if (g_ranges[i].first == (const char *)base) {
            drop_cached_locked(g_ranges[i].first, g_ranges[i].first + g_ranges[i].second);
            g_ranges.erase(g_ranges.begin() + i);
            return TB200_OK;
        }
    tb200_set_error("unregister_weights: range not registered");
    return TB200_ERR_INVALID;
}

size_t tb200_weights_cache_bytes(void) {
    std::lock_guard<std::mutex> lk(g_cache_mu);
    size_t total = 0;
    for (auto &kv : g_cache) total += kv.second->bytes;
    return total;
}

bool llamafile_sgemm(long m, long n, long k, const void *A, long lda, const void *B, long ldb, void *C, long ldc, int ith, int nth,
                     int Atype, int Btype, int Ctype) {
    // the reference asserts these (tinyblas_cpu_sgemm.inc:277-284); a violated precondition is "not handled"
    if (m < 0 || n < 0 || k < 0 || lda < k || ldb < k || ldc < m || nth <= 0 || ith < 0 || ith >= nth) return false;
    if (!tb200_serviced(Atype, Btype, Ctype, n)) return false;
    if (g_disabled.load(std::memory_order_acquire)) return false;   // same answer on every thread (see g_disabled)
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return false;
    if (ith != 0) return true;   // thread 0 owns the whole job; see the threading contract in the header
    if (m == 0 || n == 0) return true;
    DeviceLock lock(c->device);
    const long ke = k * blck(Atype);
    // iqk (Q4_0, Q4_K) takes rows ggml_row_size apart and ignores lda/ldb (iqk_mul_mat.inc:191-199); tinyBLAS honours them
    const bool iqk = Atype == TB200_TYPE_Q4_0 || Atype == TB200_TYPE_Q5_0 || Atype == TB200_TYPE_Q4_1 || tb200_is_kquant(Atype);
    const size_t a_stride = iqk ? tb200_gguf_row_bytes(Atype, ke) : (size_t)lda * tsize(Atype);
    const long ldb_eff = iqk ? k : ldb;
    WeightsRef wr;
    wr.w = cached_weights(A, Atype, m, ke, a_stride, &wr.transient);
    tb200_weights *w = wr.w;
    if (!w) return runtime_failure("llamafile_sgemm (weight upload)", nth);
    const size_t b_bytes = ((size_t)(n - 1) * ldb_eff + k) * tsize(Btype), c_bytes = (size_t)n * m * 4;
    if (n <= TB200_GEMV_MAX_TOKENS && lane_enabled()) {
        // decode lane (see HostLane)
        const bool tr = trace_enabled();
        const double t0 = tr ? now_us() : 0;
        HostLane &L = g_lane[c->device];
        if (!lane_init(L)) return runtime_failure("llamafile_sgemm (lane init)", nth);
        LaneKey key = {w, n, Btype, pinned_alias(B), nullptr, ldb_eff, ldc, nullptr};
        if (!key.dB || (Atype == TB200_TYPE_F16 && ((uintptr_t)key.dB & 15))) {   // the fp16/f32 stage copy moves 16-byte words
            if (!lane_reserve_host(L, &L.in, &L.in_cap, b_bytes)) return runtime_failure("llamafile_sgemm (staging)", nth);
            memcpy(L.in, B, b_bytes);
            key.dB = L.in;
        }
        void *pinC = pinned_alias(C);
        if (c_bytes <= lane_direct_bytes()) {   // kernels store C over PCIe
            key.dC = pinC;
            if (!key.dC) {
                if (!lane_reserve_host(L, &L.out, &L.out_cap, c_bytes)) return runtime_failure("llamafile_sgemm (staging)", nth);
                key.dC = L.out, key.ldo = m;
            }
        } else {                             // HBM, then one D2H transfer into pinned memory
            if (!lane_reserve_dev(L, &L.dout, &L.dout_cap, c_bytes)) return runtime_failure("llamafile_sgemm (staging)", nth);
            key.dC = L.dout, key.ldo = m;
            key.h_dst = (pinC && (ldc == m || n == 1)) ? (void *)C : nullptr;
            if (!key.h_dst) {
                if (!lane_reserve_host(L, &L.out, &L.out_cap, c_bytes)) return runtime_failure("llamafile_sgemm (staging)", nth);
                key.h_dst = L.out;
            }
        }
        const size_t act_bytes = Atype == TB200_TYPE_F16 ? up256(b_bytes) : up256((size_t)n * ke) + up256((size_t)n * (ke / QK) * 8);
        if (!lane_reserve_dev(L, &L.dact, &L.dact_cap, act_bytes)) return runtime_failure("llamafile_sgemm (staging)", nth);
        const double t1 = tr ? now_us() : 0;

        const uint32_t seq = ++L.seq;
        bool ok = true;
        const bool graphs = lane_graphs_enabled() && !wr.transient;   // a transient copy is gone after this call: nothing to replay
        auto it = graphs ? L.graphs.find(key) : L.graphs.end();
        if (it != L.graphs.end() && it->second) {
            ok = cudaGraphLaunch(it->second, L.stream) == cudaSuccess;
            if (tr) ++g_trace.replays;
        } else {
            cudaStream_t user = c->stream;
            c->stream = L.stream;
            const bool capture = graphs && it != L.graphs.end();   // second sighting: kernel attributes are set, buffers sized
            if (capture) ok = cudaStreamBeginCapture(L.stream, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
            if (ok) ok = lane_issue(c, L, key, b_bytes) == TB200_OK;
            if (capture) {
                cudaGraph_t g = nullptr;
                const bool ended = cudaStreamEndCapture(L.stream, &g) == cudaSuccess && g;
                cudaGraphExec_t ex = nullptr;
                ok = ok && ended && cudaGraphInstantiate(&ex, g, 0) == cudaSuccess;
                if (g) cudaGraphDestroy(g);
                if (ok) { it->second = ex; ok = cudaGraphLaunch(ex, L.stream) == cudaSuccess; }
            } else if (ok && graphs) L.graphs[key] = nullptr;
            c->stream = user;
        }
        if (!ok) {
            cudaGetLastError();
            cudaStreamSynchronize(L.stream);
            // resynchronise the host mirror with the device counter so later calls keep working
            cudaMemcpy(&L.seq, L.dcount, 4, cudaMemcpyDeviceToHost);
            return runtime_failure("llamafile_sgemm (lane launch)", nth);
        }
        const double t2 = tr ? now_us() : 0;
        if (!lane_wait(L, seq)) return runtime_failure("llamafile_sgemm (lane wait)", nth);
        const double t3 = tr ? now_us() : 0;
        const float *src = key.h_dst ? (const float *)key.h_dst : (const float *)key.dC;
        if ((const void *)src == (const void *)L.out)
            for (long j = 0; j < n; ++j) memcpy((float *)C + j * ldc, src + j * m, (size_t)m * 4);
        if (tr) {
            LaneTrace &T = g_trace;
            T.prep += t1 - t0, T.launch += t2 - t1, T.wait += t3 - t2, T.out += now_us() - t3;
            if (++T.calls % 1024 == 0)
                fprintf(stderr, "tb200 lane: %ld calls (%ld graph replays), per call us: prep %.2f launch %.2f wait %.2f copy-out %.2f\n", T.calls,
                        T.replays, T.prep / T.calls, T.launch / T.calls, T.wait / T.calls, T.out / T.calls);
        }
        return true;
    }
    uint8_t *io = (uint8_t *)io_buffer(c, up256(b_bytes) + c_bytes);
    if (!io) return runtime_failure("llamafile_sgemm (io buffer)", nth);
    float *dC = (float *)(io + up256(b_bytes));
    bool ok = cudaMemcpyAsync(io, B, b_bytes, cudaMemcpyHostToDevice, c->stream) == cudaSuccess;
    ok = ok && tb200_sgemm_dev(w, n, io, ldb_eff, Btype, dC, m) == TB200_OK;   // ldb: elements for F16/F32, blocks for quants
    ok = ok && cudaMemcpy2DAsync(C, (size_t)ldc * 4, dC, (size_t)m * 4, (size_t)m * 4, (size_t)n, cudaMemcpyDeviceToHost, c->stream) == cudaSuccess;
    ok = cudaStreamSynchronize(c->stream) == cudaSuccess && ok;
    if (!ok) { cudaGetLastError(); return runtime_failure("llamafile_sgemm", nth); }
    return true;
}

// llamafile/tinyblas_cpu_mixmul.inc:97-111 + tinyblas_cpu_mixmul_amd_avx.cpp:8-18 (arena arithmetic only)
size_t llamafile_mixmul_needs(const struct ggml_tensor *weights_, const struct ggml_tensor *thought_, const struct ggml_tensor *plan_) {
    const tb200_ggml_tensor_view *w = (const tb200_ggml_tensor_view *)weights_, *x = (const tb200_ggml_tensor_view *)thought_,
                                 *p = (const tb200_ggml_tensor_view *)plan_;
    const size_t cols = w->ne[0], experts = w->ne[2], thinkers = p->ne[0], tasks = x->ne[1], tokens = x->ne[2];
    const size_t ldq = (cols * 2 + 63) & ~(size_t)63;
    const size_t wsize = 0x7ffff000, wdata = 0x1000, wdata_aligned = (wdata + 4095) & ~(size_t)4095;
    size_t allocated = 0;
    auto alloc = [&](size_t align, size_t need) {
        const size_t base = (allocated + align - 1) & ~(align - 1), toto = base + need;
        if (toto >= allocated && toto <= wsize) { allocated = toto; return true; }
        return false;
    };
    if (!alloc(4096, tokens * tasks * ldq) || !alloc(64, 8 * experts * tokens * thinkers) || !alloc(64, 8 * experts * tokens * thinkers) ||
        !alloc(8, 8 * experts))
        return 0;
    return (wdata_aligned - wdata) + allocated;
}

// A strided host tensor [inner bytes] x [d1 rows, nb1 apart] x [d2 slabs, nb2 apart] <-> a dense device block, with as few
// transfers as the strides allow: ONE 2-D copy when the slabs continue the row pitch (nb2 == d1*nb1, the layout ggml
// gives contiguous and row-padded tensors), else one 2-D copy per slab.
static bool copy_3d(void *dev, const void *host, size_t inner, long d1, size_t nb1, long d2, size_t nb2, bool to_device, cudaStream_t st) {
    const cudaMemcpyKind kind = to_device ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost;
    auto copy2d = [&](char *d, const char *h, long rows) {
        return (to_device ? cudaMemcpy2DAsync(d, inner, h, nb1, inner, (size_t)rows, kind, st)
                          : cudaMemcpy2DAsync((void *)h, nb1, d, inner, inner, (size_t)rows, kind, st)) == cudaSuccess;
    };
    if (d1 == 1 && d2 > 1) return copy_3d(dev, host, inner, d2, nb2, 1, 0, to_device, st);
    if (d2 <= 1 || nb2 == (size_t)d1 * nb1) return copy2d((char *)dev, (const char *)host, d1 * (d2 > 0 ? d2 : 0));
    for (long s = 0; s < d2; ++s)
        if (!copy2d((char *)dev + (size_t)s * d1 * inner, (const char *)host + (size_t)s * nb2, d1)) return false;
    return true;
}

bool llamafile_mixmul(const struct ggml_compute_params *params_, const struct ggml_tensor *weights_, const struct ggml_tensor *thought_,
                      const struct ggml_tensor *plan_, struct ggml_tensor *result_) {
    const tb200_ggml_compute_params_view *params = (const tb200_ggml_compute_params_view *)params_;
    const tb200_ggml_tensor_view *w = (const tb200_ggml_tensor_view *)weights_, *x = (const tb200_ggml_tensor_view *)thought_,
                                 *pl = (const tb200_ggml_tensor_view *)plan_;
    tb200_ggml_tensor_view *out = (tb200_ggml_tensor_view *)result_;
    const long rows = w->ne[1], cols = w->ne[0], experts = w->ne[2], thinkers = pl->ne[0], tasks = x->ne[1], tokens = x->ne[2];
    // allocate_shared_memory() must succeed in the caller's wdata, as in the reference (:97-107, :390-398):
    // offsets are measured from the 4096-aligned start and compared with params->wsize
    {
        const size_t need = llamafile_mixmul_needs(weights_, thought_, plan_);
        if (!need || need > params->wsize) return false;
    }
    // invariants the reference asserts (:116-135)
    if (!(tasks <= thinkers && thinkers <= experts && tokens == pl->ne[1] && rows == out->ne[0] && cols == x->ne[0] &&
          tokens == out->ne[2] && thinkers == out->ne[1] && pl->type == TB200_TYPE_I32 && params->nth > 0 && params->ith < params->nth))
        return false;
    if (pl->ne[2] != 1 || pl->ne[3] != 1 || out->ne[3] != 1 || w->ne[3] != 1 || x->ne[3] != 1) return false;
    const size_t wts = tsize(w->type);
    if (!wts || w->nb[1] % wts) return false;                                              // :138
    if (out->nb[0] != 4 || x->nb[0] != tsize(x->type) || w->nb[0] != wts) return false;    // :142-147
    if (out->type != TB200_TYPE_F32) return false;
    if (w->type != TB200_TYPE_Q4_0 && w->type != TB200_TYPE_Q8_0 && w->type != TB200_TYPE_F16) return false;
    if (x->type != TB200_TYPE_F32) return false;   // pre-converted thought (Q8_0/F16) is never produced by ggml for this op
    if (cols % (w->type == TB200_TYPE_F16 ? 16 : 32)) return false;                        // :273
    if (w->nb[2] != (size_t)rows * w->nb[1]) return false;                                 // experts must be stacked
    if (pl->nb[0] != 4) return false;
    if (g_disabled.load(std::memory_order_acquire)) return false;
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return false;
    if (params->ith != 0) return true;
    if (tokens == 0 || rows == 0) return true;
    const int nth = params->nth;
    DeviceLock lock(c->device);
    WeightsRef wr;
    wr.w = cached_weights(w->data, w->type, rows * experts, cols, w->nb[1], &wr.transient);
    tb200_weights *dw = wr.w;
    if (!dw) return runtime_failure("llamafile_mixmul (weight upload)", nth);
    const size_t x_bytes = (size_t)tokens * tasks * cols * 4, p_bytes = (size_t)tokens * thinkers * 4, o_bytes = (size_t)tokens * thinkers * rows * 4;
    uint8_t *io = (uint8_t *)io_buffer(c, up256(x_bytes) + up256(p_bytes) + o_bytes);
    if (!io) return runtime_failure("llamafile_mixmul (io buffer)", nth);
    float *dx = (float *)io;
    int32_t *dp = (int32_t *)(io + up256(x_bytes));
    float *dout = (float *)(io + up256(x_bytes) + up256(p_bytes));
    // the (possibly strided) host tensors <-> dense device blocks: one transfer each when the strides allow (copy_3d)
    bool ok = copy_3d(dx, x->data, (size_t)cols * 4, tasks, x->nb[1], tokens, x->nb[2], true, c->stream);
    ok = ok && copy_3d(dp, pl->data, (size_t)thinkers * 4, tokens, pl->nb[1], 1, 0, true, c->stream);
    bool all_valid = true;
    for (long t = 0; t < tokens && all_valid; ++t)
        for (long th = 0; th < thinkers; ++th) {
            const int32_t e = *(const int32_t *)((const char *)pl->data + t * pl->nb[1] + th * pl->nb[0]);
            if (e < 0 || e >= experts) { all_valid = false; break; }
        }
    // rows naming no expert stay untouched in the reference: carry the caller's bytes through the device copy
    if (ok && !all_valid) ok = copy_3d(dout, out->data, (size_t)rows * 4, thinkers, out->nb[1], tokens, out->nb[2], true, c->stream);
    if (ok) ok = tb200_mixmul_dev(dw, (int)experts, dx, (int)tasks, dp, (int)thinkers, tokens, dout) == TB200_OK;
    ok = ok && copy_3d(dout, out->data, (size_t)rows * 4, thinkers, out->nb[1], tokens, out->nb[2], false, c->stream);
    ok = cudaStreamSynchronize(c->stream) == cudaSuccess && ok;
    if (!ok) { cudaGetLastError(); return runtime_failure("llamafile_mixmul", nth); }
    return true;
}

// iqk_mul_mat_moe (llamafile/iqk_mul_mat.inc:204-221): A = one expert's Nx rows, B = all quantized rows
// [ne12][ne11], C row for mapping y = C + i1*nb1 + i2*nb2, B row = (i1 % ne11) + i2*ne11.
bool llamafile_mixmul_iqk(long Nx, long Ny, long ne00, int ne11, int typeA, const void *A, const void *B, float *C, long nb1, long nb2,
                          const void *vrow_mapping, int ith, int nth) {
    if (!tb200_is_kquant(typeA) && typeA != TB200_TYPE_Q4_0 && typeA != TB200_TYPE_Q5_0 && typeA != TB200_TYPE_Q4_1) return false;   // the i-quants below 4 bits and Q5_1 / IQ4_NL are outside this module
    if (!vrow_mapping || Nx < 0 || Ny < 0 || ne11 <= 0 || nth <= 0 || ith < 0 || ith >= nth || ne00 % blck(typeA)) return false;
    if (g_disabled.load(std::memory_order_acquire)) return false;
    tb200_ctx *c = tb200_get_ctx();
    if (!c) return false;
    if (ith != 0) return true;
    if (Nx == 0 || Ny == 0) return true;
    DeviceLock lock(c->device);
    const int Btype = tb200_is_kquant(typeA) ? TB200_TYPE_Q8_K : (typeA == TB200_TYPE_Q4_1 ? TB200_TYPE_Q8_1 : TB200_TYPE_Q8_0);
    WeightsRef wr;
    wr.w = cached_weights(A, typeA, Nx, ne00, 0, &wr.transient);
    tb200_weights *w = wr.w;
    if (!w) return runtime_failure("llamafile_mixmul_iqk (weight upload)", nth);
    const int32_t *rm = (const int32_t *)vrow_mapping;
    const size_t row_b = tb200_gguf_row_bytes(Btype, ne00), in_b = up256(row_b * Ny), out_b = (size_t)Ny * Nx * 4;
    // The mapped B rows are gathered on the host into ONE pinned staging block (plain memcpy, no driver call per row), go
    // to the device in one transfer, one dense sgemm runs, and the result comes back in one transfer and is scattered to
    // the caller's rows the same way. (The first version issued a pageable cudaMemcpyAsync per row each way: > 1000 driver
    // calls around a 300 us kernel at Ny = 512.)
    HostLane &L = g_lane[c->device];
    if (!lane_reserve_host(L, &L.in, &L.in_cap, in_b) || !lane_reserve_host(L, &L.out, &L.out_cap, out_b))
        return runtime_failure("llamafile_mixmul_iqk (staging)", nth);
    uint8_t *io = (uint8_t *)io_buffer(c, in_b + out_b);
    if (!io) return runtime_failure("llamafile_mixmul_iqk (io buffer)", nth);
    float *dC = (float *)(io + in_b);
    for (long y = 0; y < Ny; ++y) {
        const long brow = (long)(rm[2 * y] % ne11) + (long)rm[2 * y + 1] * ne11;
        memcpy(L.in + y * row_b, (const char *)B + brow * row_b, row_b);
    }
    bool ok = cudaMemcpyAsync(io, L.in, row_b * Ny, cudaMemcpyHostToDevice, c->stream) == cudaSuccess;
    ok = ok && tb200_sgemm_dev(w, Ny, io, ne00 / blck(typeA), Btype, dC, Nx) == TB200_OK;
    ok = ok && cudaMemcpyAsync(L.out, dC, out_b, cudaMemcpyDeviceToHost, c->stream) == cudaSuccess;
    ok = cudaStreamSynchronize(c->stream) == cudaSuccess && ok;
    if (!ok) { cudaGetLastError(); return runtime_failure("llamafile_mixmul_iqk", nth); }
    for (long y = 0; y < Ny; ++y)
        memcpy(C + (long)rm[2 * y] * (nb1 / 4) + (long)rm[2 * y + 1] * (nb2 / 4), (const float *)L.out + y * Nx, (size_t)Nx * 4);
    return true;
}

static bool quantize_row_host(const float *x, void *y, int64_t k, int type) {
    tb200_ctx *c = tb200_get_ctx();
    if (!c || k < 0) return false;
    if (k == 0) return true;
    DeviceLock lock(c->device);
    const size_t out_b = type == TB200_TYPE_F16 ? (size_t)k * 2 : tb200_gguf_row_bytes(type, k);
    if (!out_b) return false;
    uint8_t *io = (uint8_t *)io_buffer(c, up256((size_t)k * 4) + out_b);
    if (!io) return false;
    uint8_t *dy = io + up256((size_t)k * 4);
    if (cudaMemcpyAsync(io, x, (size_t)k * 4, cudaMemcpyHostToDevice, c->stream) != cudaSuccess) { cudaGetLastError(); return false; }
    int rc;
    if (type == TB200_TYPE_Q8_0) rc = tb200_launch_quantize_q8_0(c, (const float *)io, k, 1, k, dy, nullptr, nullptr);
    else if (type == TB200_TYPE_Q8_K) rc = tb200_launch_quantize_q8_K(c, (const float *)io, k, 1, k, dy, nullptr, nullptr);
    else rc = tb200_launch_f32_to_f16(c, (const float *)io, k, 1, k, (__half *)dy);
    if (rc) return false;
    if (cudaMemcpyAsync(y, dy, out_b, cudaMemcpyDeviceToHost, c->stream) != cudaSuccess) { cudaGetLastError(); return false; }
    return cudaStreamSynchronize(c->stream) == cudaSuccess;
}
bool tb200_quantize_row_q8_0(const float *x, void *y, int64_t k) { return k % QK == 0 && quantize_row_host(x, y, k, TB200_TYPE_Q8_0); }
bool tb200_quantize_row_q8_K(const float *x, void *y, int64_t k) { return k % QK_K == 0 && quantize_row_host(x, y, k, TB200_TYPE_Q8_K); }
bool tb200_fp32_to_fp16_row(const float *x, uint16_t *y, int64_t k) { return quantize_row_host(x, y, k, TB200_TYPE_F16); }

}  // extern "C"
