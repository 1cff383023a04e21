// tinyBLAS-B200 as a ggml backend (include/tb200_ggml_backend.h): buffer type, buffers, and a backend whose
// graph_compute runs GGML_OP_MUL_MAT / GGML_OP_MUL_MAT_ID nodes on device-resident tensors without returning to the host
// between nodes. Mirrors the reference's GPU backend object by object: interface table llama.cpp/ggml/src/ggml-cuda.cu:2976-2997,
// graph_compute :2467, supports_op :2725, buffer type :548-640, host buffer type :1040-1100, loader handshake
// llamafile/cuda.c:701-753 + ggml_cuda_link (llama.cpp.patches/patches/ggml-cuda.cu.patch:468-520). Host-side C++ only; every
// kernel is the module's own (tb200_*.cu). Nothing here links against ggml: the host's helpers arrive through the
// ggml_backend_api table, as in llamafile.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <map>
#include <unordered_map>
#include <mutex>
#include <vector>

#include "../../include/tb200_ggml_backend.h"
#include "tb200_common.cuh"

size_t tb200_gguf_row_bytes(int type, long k);

namespace {

typedef tb200_ggml_tensor_view T;
inline const T *tv(const struct ggml_tensor *t) { return reinterpret_cast<const T *>(t); }
inline T *tv(struct ggml_tensor *t) { return reinterpret_cast<T *>(t); }

tb200_ggml_backend_api g_api;
bool g_linked = false;
constexpr int kMaxDev = 16;
const uint8_t g_guid[16] = {0x7b, 0x20, 0x0b, 0x1a, 0x5b, 0x10, 0x0a, 0x4c, 0x9d, 0x11, 0x62, 0x32, 0x30, 0x30, 0x42, 0x54};

void backend_log(const char *msg) {
    if (g_api.FLAG_log_disable && *g_api.FLAG_log_disable) return;
    if (g_api.write) g_api.write(2, msg, (long)strlen(msg));
    else fputs(msg, stderr);
}

bool weight_type_ok(int type) {
    return type == TB200_TYPE_F16 || type == TB200_TYPE_Q4_0 || type == TB200_TYPE_Q8_0 || type == TB200_TYPE_Q4_K || type == TB200_TYPE_Q6_K || tb200_is_g8(type);
}
long type_blck(int type) { return type == TB200_TYPE_F16 ? 1 : (tb200_is_kquant(type) ? QK_K : QK); }

// ------------------------------------------------------------------------------------------------ buffers
struct Resident { tb200_weights *w; int type; long k, rows; size_t stride; };
// bumped whenever a resident (repacked) copy is freed: cached decode plans hold such copies and are rebuilt when it moves
std::atomic<uint64_t> g_resident_epoch{1};
struct BufferCtx {
    int device;
    void *base;
    size_t size;
    std::mutex mu;
    std::map<const void *, Resident> resident;   // weight tensors repacked for the kernels, keyed by the tensor's data pointer
    void drop(const void *lo, const void *hi) {  // raw bytes in [lo, hi) changed: the repacked copies made from them are stale
        std::lock_guard<std::mutex> lk(mu);
        for (auto it = resident.begin(); it != resident.end();) {
            const char *a = (const char *)it->first, *b = a + it->second.stride * it->second.rows;
            if (!lo || (a < (const char *)hi && b > (const char *)lo)) {
                cudaDeviceSynchronize();                     // a plan or kernel still in flight may be reading it
                tb200_weights_free(it->second.w);
                it = resident.erase(it);
                g_resident_epoch.fetch_add(1);
            }
            else ++it;
        }
    }
};

struct BuftCtx { int device; char name[16]; };
tb200_ggml_backend_buffer_type g_buft[kMaxDev];
BuftCtx g_buft_ctx[kMaxDev];
tb200_ggml_backend_buffer_type g_host_buft;

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

const char *buffer_get_name(tb200_buffer_t b) { return g_buft_ctx[((BufferCtx *)b->context)->device].name; }
void buffer_free(tb200_buffer_t b) {
    BufferCtx *c = (BufferCtx *)b->context;
    DeviceGuard g(c->device);
    cudaDeviceSynchronize();
    c->drop(nullptr, nullptr);
    if (c->base) cudaFree(c->base);
    delete c;
}
void *buffer_get_base(tb200_buffer_t b) { return ((BufferCtx *)b->context)->base; }
void buffer_init_tensor(tb200_buffer_t, struct ggml_tensor *) {}
void buffer_set_tensor(tb200_buffer_t b, struct ggml_tensor *t, const void *data, size_t offset, size_t size) {
    BufferCtx *c = (BufferCtx *)b->context;
    DeviceGuard g(c->device);
    char *dst = (char *)tv(t)->data + offset;
    c->drop(dst, dst + size);
    cudaMemcpyAsync(dst, data, size, cudaMemcpyDefault, cudaStreamPerThread);   // `data` is normally host memory; a device pointer works too (UVA)
    cudaStreamSynchronize(cudaStreamPerThread);
}
void buffer_get_tensor(tb200_buffer_t b, const struct ggml_tensor *t, void *data, size_t offset, size_t size) {
    BufferCtx *c = (BufferCtx *)b->context;
    DeviceGuard g(c->device);
    cudaMemcpyAsync(data, (const char *)tv(t)->data + offset, size, cudaMemcpyDefault, cudaStreamPerThread);
    cudaStreamSynchronize(cudaStreamPerThread);
}
bool buffer_cpy_tensor(tb200_buffer_t b, const struct ggml_tensor *src, struct ggml_tensor *dst) {
    const tb200_buffer_t sb = (tb200_buffer_t)tv(src)->buffer_;
    if (!sb || sb->iface.get_name != buffer_get_name || !g_api.ggml_nbytes) return false;   // only between this module's buffers
    BufferCtx *c = (BufferCtx *)b->context, *sc = (BufferCtx *)sb->context;
    const size_t n = g_api.ggml_nbytes(src);
    DeviceGuard g(c->device);
    c->drop(tv(dst)->data, (char *)tv(dst)->data + n);
    if (sc->device == c->device) cudaMemcpyAsync(tv(dst)->data, tv(src)->data, n, cudaMemcpyDeviceToDevice, cudaStreamPerThread);
    else cudaMemcpyPeerAsync(tv(dst)->data, c->device, tv(src)->data, sc->device, n, cudaStreamPerThread);
    cudaStreamSynchronize(cudaStreamPerThread);
    return true;
}
void buffer_clear(tb200_buffer_t b, uint8_t value) {
    BufferCtx *c = (BufferCtx *)b->context;
    DeviceGuard g(c->device);
    c->drop(nullptr, nullptr);
    cudaDeviceSynchronize();
    cudaMemset(c->base, value, c->size);
    cudaDeviceSynchronize();
}
void buffer_reset(tb200_buffer_t b) { ((BufferCtx *)b->context)->drop(nullptr, nullptr); }

const tb200_ggml_backend_buffer_i kBufferIface = {buffer_get_name, buffer_free, buffer_get_base, buffer_init_tensor, buffer_set_tensor,
                                                  buffer_get_tensor, buffer_cpy_tensor, buffer_clear, buffer_reset};

const char *buft_get_name(tb200_buft_t t) { return ((BuftCtx *)t->context)->name; }
tb200_buffer_t buft_alloc_buffer(tb200_buft_t t, size_t size) {
    BuftCtx *bc = (BuftCtx *)t->context;
    DeviceGuard g(bc->device);
    BufferCtx *c = new BufferCtx();
    c->device = bc->device, c->size = size > 1 ? size : 1, c->base = nullptr;
    if (cudaMalloc(&c->base, c->size) != cudaSuccess) {
        cudaGetLastError();
        char msg[160];
        snprintf(msg, sizeof msg, "tb200_backend: allocating %.2f MiB on device %d failed\n", size / 1048576.0, bc->device);
        backend_log(msg);
        delete c;
        return nullptr;
    }
    return g_api.ggml_backend_buffer_init(t, kBufferIface, c, size);
}
size_t buft_get_alignment(tb200_buft_t) { return 256; }   // every tensor 16-byte aligned for the kernels' vector loads and TMA (cudaMalloc gives 256)
bool buft_is_host(tb200_buft_t) { return false; }

// pinned host buffers: the host's CPU buffer around cudaMallocHost memory (ggml-cuda.cu:1040-1100)
const char *host_buft_get_name(tb200_buft_t) { return "TB200_Host"; }
void host_buffer_free(tb200_buffer_t b) { cudaFreeHost(b->context); }
tb200_buffer_t host_buft_alloc_buffer(tb200_buft_t t, size_t size) {
    void *p = nullptr;
    if (cudaMallocHost(&p, size > 1 ? size : 1) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    tb200_buffer_t b = g_api.ggml_backend_cpu_buffer_from_ptr(p, size);
    b->buft = t;
    b->iface.free_buffer = host_buffer_free;   // the CPU buffer's context is the pointer itself
    return b;
}
size_t host_buft_get_alignment(tb200_buft_t) { return 64; }
bool host_buft_is_host(tb200_buft_t) { return true; }

// ------------------------------------------------------------------------------------------------ backend
struct PlanKey {
    std::vector<const void *> ptrs;   // (output data, weights data, input data, op | first parameter word, row length) per node
    bool operator<(const PlanKey &o) const { return ptrs < o.ptrs; }
};
struct BackendCtx {
    int device;
    cudaStream_t stream;
    struct PlanEntry { tb200_plan *plan; int groups; uint64_t epoch; int end, matmuls; };   // end: first node after the run the plan covers
    std::map<PlanKey, PlanEntry> plans;
    void *ids_scratch = nullptr; size_t ids_cap = 0;
    int last_launches = 0, last_plan_nodes = 0;
    char name[16];
};

const char *backend_get_name(tb200_backend_t b) { return ((BackendCtx *)b->context)->name; }
void backend_free(tb200_backend_t b) {
    BackendCtx *c = (BackendCtx *)b->context;
    DeviceGuard g(c->device);
    cudaStreamSynchronize(c->stream);
    for (auto &kv : c->plans) tb200_plan_free(kv.second.plan);
    if (c->ids_scratch) cudaFree(c->ids_scratch);
    cudaStreamDestroy(c->stream);
    delete c;
    delete b;
}
tb200_buft_t backend_default_buft(tb200_backend_t b) { return tb200_backend_buffer_type(((BackendCtx *)b->context)->device); }
void backend_set_tensor_async(tb200_backend_t b, struct ggml_tensor *t, const void *data, size_t offset, size_t size) {
    BackendCtx *c = (BackendCtx *)b->context;
    DeviceGuard g(c->device);
    tb200_buffer_t buf = (tb200_buffer_t)(tv(t)->view_src_ ? tv((struct ggml_tensor *)tv(t)->view_src_)->buffer_ : tv(t)->buffer_);
    char *dst = (char *)tv(t)->data + offset;
    if (buf && buf->iface.get_name == buffer_get_name) ((BufferCtx *)buf->context)->drop(dst, dst + size);
    cudaMemcpyAsync(dst, data, size, cudaMemcpyDefault, c->stream);
}
void backend_get_tensor_async(tb200_backend_t b, const struct ggml_tensor *t, void *data, size_t offset, size_t size) {
    BackendCtx *c = (BackendCtx *)b->context;
    DeviceGuard g(c->device);
    cudaMemcpyAsync(data, (const char *)tv(t)->data + offset, size, cudaMemcpyDefault, c->stream);
}
void backend_synchronize(tb200_backend_t b) {
    BackendCtx *c = (BackendCtx *)b->context;
    DeviceGuard g(c->device);
    cudaStreamSynchronize(c->stream);
}

// weights of a MUL_MAT / MUL_MAT_ID node: repacked on the device the first time a graph uses them, then resident
tb200_weights *resident_weights(const T *a, long rows_total) {
    tb200_buffer_t buf = (tb200_buffer_t)(a->view_src_ ? tv((const struct ggml_tensor *)a->view_src_)->buffer_ : a->buffer_);
    if (!buf || buf->iface.get_name != buffer_get_name) { tb200_set_error("backend: weights are not in a TB200 buffer"); return nullptr; }
    BufferCtx *bc = (BufferCtx *)buf->context;
    std::lock_guard<std::mutex> lk(bc->mu);
    auto it = bc->resident.find(a->data);
    if (it != bc->resident.end() && it->second.type == a->type && it->second.k == a->ne[0] && it->second.rows == rows_total && it->second.stride == a->nb[1])
        return it->second.w;
    if (it != bc->resident.end()) { cudaDeviceSynchronize(); tb200_weights_free(it->second.w); bc->resident.erase(it); g_resident_epoch.fetch_add(1); }
    tb200_weights *w = tb200_weights_from_device(a->data, a->type, rows_total, a->ne[0], a->nb[1]);
    if (w) bc->resident[a->data] = Resident{w, a->type, (long)a->ne[0], rows_total, a->nb[1]};
    return w;
}

bool mul_mat_supported(const T *op) {
    const T *a = tv((const struct ggml_tensor *)op->src_[0]), *b = tv((const struct ggml_tensor *)op->src_[1]);
    if (!a || !b || !weight_type_ok(a->type) || b->type != TB200_TYPE_F32 || op->type != TB200_TYPE_F32) return false;
    if (a->ne[2] != 1 || a->ne[3] != 1 || b->ne[2] != 1 || b->ne[3] != 1) return false;          // no batched / broadcast matmuls (attention stays elsewhere)
    if (a->ne[0] != b->ne[0] || a->ne[0] % type_blck(a->type) || a->ne[0] < 1 || a->ne[1] < 1) return false;
    const size_t rb = tb200_gguf_row_bytes(a->type, a->ne[0]);
    if (a->nb[0] != tb200_gguf_row_bytes(a->type, type_blck(a->type)) || a->nb[1] < rb) return false;   // rows contiguous inside, any row pitch
    if (b->nb[0] != 4 || b->nb[1] % 4 || b->nb[1] < (size_t)b->ne[0] * 4) return false;
    if (op->nb[0] != 4 || op->nb[1] % 4 || op->nb[1] < (size_t)op->ne[0] * 4 || op->ne[0] != a->ne[1] || op->ne[1] != b->ne[1]) return false;
    return true;
}
bool mul_mat_id_supported(const T *op) {
    const T *as = tv((const struct ggml_tensor *)op->src_[0]), *b = tv((const struct ggml_tensor *)op->src_[1]), *ids = tv((const struct ggml_tensor *)op->src_[2]);
    if (!as || !b || !ids) return false;
    if (as->type != TB200_TYPE_Q4_0 && as->type != TB200_TYPE_Q8_0 && as->type != TB200_TYPE_F16) return false;   // what llamafile_mixmul serves
    if (b->type != TB200_TYPE_F32 || op->type != TB200_TYPE_F32 || ids->type != TB200_TYPE_I32) return false;
    if (as->ne[3] != 1 || b->ne[3] != 1 || ids->ne[2] != 1 || ids->ne[3] != 1 || op->ne[3] != 1) return false;
    const long K = as->ne[0], rows = as->ne[1], experts = as->ne[2], tasks = b->ne[1], tokens = b->ne[2], thinkers = ids->ne[0];
    if (K != b->ne[0] || K % (as->type == TB200_TYPE_F16 ? 16 : 32) || ids->ne[1] != tokens || (tasks != 1 && tasks != thinkers) || thinkers > experts) return false;
    if (op->ne[0] != rows || op->ne[1] != thinkers || op->ne[2] != tokens) return false;
    const size_t rb = tb200_gguf_row_bytes(as->type, K);
    if (as->nb[1] != rb || as->nb[2] != rb * rows) return false;                                   // experts stacked densely
    if (b->nb[0] != 4 || b->nb[1] != (size_t)K * 4 || b->nb[2] != (size_t)K * 4 * tasks) return false;
    if (ids->nb[0] != 4 || ids->nb[1] % 4) return false;
    if (op->nb[0] != 4 || op->nb[1] != (size_t)rows * 4 || op->nb[2] != (size_t)rows * 4 * thinkers) return false;
    return true;
}
bool is_view_op(int op) {
    return op == TB200_GGML_OP_NONE || op == TB200_GGML_OP_RESHAPE || op == TB200_GGML_OP_VIEW || op == TB200_GGML_OP_PERMUTE || op == TB200_GGML_OP_TRANSPOSE;
}
// the matmul's neighbours (llm_build_norm, llm_build_ffn: llama.cpp/src/llama.cpp:8099-8225): fp32 matrices [ne0 x ne1] with dense rows;
// MUL / ADD take a second operand of the same shape or one row broadcast over the tokens (the norm weight)
bool f32_rows(const T *t) { return t && t->type == TB200_TYPE_F32 && t->ne[2] == 1 && t->ne[3] == 1 && t->nb[0] == 4 && t->nb[1] % 4 == 0 && t->nb[1] >= (size_t)t->ne[0] * 4; }
bool ew_supported(const T *op) {
    const T *a = tv((const struct ggml_tensor *)op->src_[0]), *b = tv((const struct ggml_tensor *)op->src_[1]);
    if (!f32_rows(op) || !f32_rows(a) || a->ne[0] != op->ne[0] || a->ne[1] != op->ne[1]) return false;
    switch (op->op_) {
    case TB200_GGML_OP_RMS_NORM: return true;
    case TB200_GGML_OP_UNARY: return op->op_params_[0] == TB200_GGML_UNARY_OP_SILU;
    case TB200_GGML_OP_MUL: case TB200_GGML_OP_ADD: return f32_rows(b) && b->ne[0] == a->ne[0] && (b->ne[1] == a->ne[1] || b->ne[1] == 1);
    default: return false;
    }
}
bool is_ew_op(int op) { return op == TB200_GGML_OP_RMS_NORM || op == TB200_GGML_OP_UNARY || op == TB200_GGML_OP_MUL || op == TB200_GGML_OP_ADD; }
bool backend_supports_op(tb200_backend_t, const struct ggml_tensor *op_) {
    const T *op = tv(op_);
    if (is_view_op(op->op_)) return true;
    if (op->op_ == TB200_GGML_OP_MUL_MAT) return mul_mat_supported(op);
    if (op->op_ == TB200_GGML_OP_MUL_MAT_ID) return mul_mat_id_supported(op);
    if (is_ew_op(op->op_)) return ew_supported(op);
    return false;
}
bool backend_supports_buft(tb200_backend_t b, tb200_buft_t t) {
    return t && t->iface.get_name == buft_get_name && ((BuftCtx *)t->context)->device == ((BackendCtx *)b->context)->device;
}
bool backend_offload_op(tb200_backend_t, const struct ggml_tensor *) { return false; }

// a node the decode plan can take: one token, Q4_0 or Q4_K weights, K a multiple of 256, dense vectors. Returns the weight type
// (a plan holds one type: a run of plannable nodes ends where the type changes), 0 when the node cannot be planned.
int plan_node(const T *op) {
    if (op->op_ != TB200_GGML_OP_MUL_MAT || !mul_mat_supported(op)) return 0;
    const T *a = tv((const struct ggml_tensor *)op->src_[0]), *b = tv((const struct ggml_tensor *)op->src_[1]);
    const bool ok = (a->type == TB200_TYPE_Q4_0 || a->type == TB200_TYPE_Q4_K) && b->ne[1] == 1 && a->ne[0] % 256 == 0 && ((uintptr_t)b->data & 15) == 0;
    return ok ? a->type : 0;
}

int run_mul_mat(BackendCtx *, const T *op) {
    const T *a = tv((const struct ggml_tensor *)op->src_[0]), *b = tv((const struct ggml_tensor *)op->src_[1]);
    tb200_weights *w = resident_weights(a, a->ne[1]);
    if (!w) return TB200_ERR_CUDA;
    return tb200_mul_mat(w, (const float *)b->data, (long)(b->nb[1] / 4), b->ne[1], (float *)op->data, (long)(op->nb[1] / 4));
}

int run_mul_mat_id(BackendCtx *c, const T *op) {
    const T *as = tv((const struct ggml_tensor *)op->src_[0]), *b = tv((const struct ggml_tensor *)op->src_[1]), *ids = tv((const struct ggml_tensor *)op->src_[2]);
    const long rows = as->ne[1], experts = as->ne[2], tasks = b->ne[1], tokens = b->ne[2], thinkers = ids->ne[0];
    tb200_weights *w = resident_weights(as, rows * experts);
    if (!w) return TB200_ERR_CUDA;
    const int32_t *plan = (const int32_t *)ids->data;
    if (ids->nb[1] != (size_t)thinkers * 4) {   // ids is usually a view of the top-k result with the row pitch of n_expert: pack it (device to device)
        const size_t need = (size_t)tokens * thinkers * 4;
        if (need > c->ids_cap) {
            cudaStreamSynchronize(c->stream);
            if (c->ids_scratch) cudaFree(c->ids_scratch);
            c->ids_scratch = nullptr, c->ids_cap = 0;
            if (cudaMalloc(&c->ids_scratch, need * 2) != cudaSuccess) { cudaGetLastError(); return TB200_ERR_CUDA; }
            c->ids_cap = need * 2;
        }
        if (cudaMemcpy2DAsync(c->ids_scratch, (size_t)thinkers * 4, ids->data, ids->nb[1], (size_t)thinkers * 4, (size_t)tokens, cudaMemcpyDeviceToDevice, c->stream) != cudaSuccess) {
            cudaGetLastError();
            return TB200_ERR_CUDA;
        }
        plan = (const int32_t *)c->ids_scratch;
    }
    return tb200_mixmul_dev(w, (int)experts, (const float *)b->data, (int)tasks, plan, (int)thinkers, tokens, (float *)op->data);
}

int run_ew(const T *op) {
    const T *a = tv((const struct ggml_tensor *)op->src_[0]), *b = tv((const struct ggml_tensor *)op->src_[1]);
    const long k = op->ne[0], tokens = op->ne[1], lda = (long)(a->nb[1] / 4), ldy = (long)(op->nb[1] / 4);
    float *y = (float *)op->data;
    switch (op->op_) {
    case TB200_GGML_OP_RMS_NORM: { float eps; memcpy(&eps, op->op_params_, 4); return tb200_rms_norm_dev((const float *)a->data, lda, eps, tokens, k, y, ldy); }
    case TB200_GGML_OP_UNARY: return tb200_silu_dev((const float *)a->data, lda, tokens, k, y, ldy);
    case TB200_GGML_OP_MUL: return tb200_mul_dev((const float *)a->data, lda, (const float *)b->data, b->ne[1] == 1 && tokens > 1 ? 0 : (long)(b->nb[1] / 4), tokens, k, y, ldy);
    default: return tb200_add_dev((const float *)a->data, lda, (const float *)b->data, b->ne[1] == 1 && tokens > 1 ? 0 : (long)(b->nb[1] / 4), tokens, k, y, ldy);
    }
}

// ---- a run of plannable nodes starting at node i, with the matmul's neighbours folded into the decode plan --------------------------
// One-token RMS_NORM -> MUL(norm weight) feeding matmuls becomes the TB200_PLAN_RMSNORM prologue of their group, UNARY(SILU) ->
// MUL(up) the TB200_PLAN_SWIGLU prologue, an ADD of a matmul's result with another tensor its residual epilogue. The intermediates
// are then never written, so a pattern is only folded when NOTHING else reads them (every consumer in the graph is the folding node,
// and they are not graph outputs); otherwise the run ends before the pattern and those nodes run as kernels of their own.
struct FusedRun {
    std::vector<tb200_plan_group> groups;
    int end = 0, matmuls = 0, folded = 0;
};
typedef std::unordered_map<const void *, std::vector<int>> UseMap;   // tensor -> indices of the nodes that read it

bool build_fused_run(const tb200_ggml_cgraph_view *cg, int i, int limit, const UseMap &uses, FusedRun &R, int &cut) {
    struct Virt { int kind; const float *x, *y; float eps; int start, fused_uses; };   // kind: 1 NORM, 2 NORM*weight, 3 SILU, 4 SILU*up
    std::unordered_map<const void *, Virt> virt;
    std::unordered_map<const void *, std::pair<int, int>> slot;   // matmul (or its ADD) node -> (group, matrix)
    std::vector<const void *> group_in;                            // identity of each group's input (tensor), for joining
    R.groups.clear(), R.matmuls = R.folded = 0;
    int run_type = 0, j = i;
    auto n_uses = [&](const void *t) { auto it = uses.find(t); return it == uses.end() ? 0 : (int)it->second.size(); };
    auto is_output = [&](const T *t) { return (t->flags_ & TB200_GGML_TENSOR_FLAG_OUTPUT) != 0; };
    for (; j < limit; ++j) {
        const struct ggml_tensor *nt = cg->nodes[j];
        const T *n = tv(nt);
        if (is_view_op(n->op_)) {
            if (n->src_[0] && virt.count(n->src_[0])) break;      // a view of something we would not write
            continue;
        }
        const T *a = tv((const struct ggml_tensor *)n->src_[0]), *b = tv((const struct ggml_tensor *)n->src_[1]);
        if (n->op_ == TB200_GGML_OP_MUL_MAT) {
            const int t = plan_node(n);
            if (!t || (run_type && t != run_type)) break;
            tb200_plan_group pg;
            memset(&pg, 0, sizeof pg);
            auto v = virt.find(n->src_[1]);
            if (v != virt.end()) {
                if (v->second.kind == 2) pg.prologue = TB200_PLAN_RMSNORM, pg.x = v->second.x, pg.norm_weight = v->second.y, pg.eps = v->second.eps;
                else if (v->second.kind == 4) pg.prologue = TB200_PLAN_SWIGLU, pg.x = v->second.x, pg.x2 = v->second.y;
                else { cut = v->second.start; return false; }
                if (((uintptr_t)pg.x | (uintptr_t)pg.x2 | (uintptr_t)pg.norm_weight) & 15) { cut = v->second.start; return false; }
                v->second.fused_uses++;
            } else pg.x = (const float *)b->data, pg.prologue = TB200_PLAN_QUANT;
            tb200_weights *w = resident_weights(a, a->ne[1]);
            if (!w) { cut = j; return false; }
            run_type = t;
            // joins the previous group when it reads the same input (and therefore none of that group's outputs)
            const bool join = !R.groups.empty() && group_in.back() == n->src_[1] && R.groups.back().n < TB200_GEMV_MAX_MATS && R.groups.back().w[0]->k == w->k;
            if (!join) R.groups.push_back(pg), group_in.push_back(n->src_[1]);
            tb200_plan_group &G = R.groups.back();
            G.w[G.n] = w, G.out[G.n] = (float *)n->data;
            slot[nt] = std::make_pair((int)R.groups.size() - 1, G.n);
            G.n++, R.matmuls++;
            continue;
        }
        if (!is_ew_op(n->op_) || !ew_supported(n) || n->ne[1] != 1) break;
        if (n->op_ == TB200_GGML_OP_RMS_NORM || n->op_ == TB200_GGML_OP_UNARY) {
            if (virt.count(n->src_[0])) break;
            Virt v = {n->op_ == TB200_GGML_OP_RMS_NORM ? 1 : 3, (const float *)a->data, nullptr, 0.f, j, 0};
            if (v.kind == 1) memcpy(&v.eps, n->op_params_, 4);
            virt[nt] = v;
            continue;
        }
        if (n->op_ == TB200_GGML_OP_MUL) {
            auto va = virt.find(n->src_[0]);
            if (va == virt.end() || virt.count(n->src_[1]) || (va->second.kind != 1 && va->second.kind != 3)) { if (va != virt.end()) { cut = va->second.start; return false; } break; }
            if (va->second.kind == 3 && b->ne[1] != 1) break;
            va->second.fused_uses++;
            virt[nt] = Virt{va->second.kind + 1, va->second.x, (const float *)b->data, va->second.eps, va->second.start, 0};
            continue;
        }
        // ADD: residual epilogue of a matmul of this run whose result nothing else reads
        if (virt.count(n->src_[0]) || virt.count(n->src_[1])) break;
        const struct ggml_tensor *mm = slot.count(n->src_[0]) ? (const struct ggml_tensor *)n->src_[0] : (slot.count(n->src_[1]) ? (const struct ggml_tensor *)n->src_[1] : nullptr);
        if (!mm) break;
        const T *other = mm == n->src_[0] ? b : a;
        const std::pair<int, int> gm = slot[mm];
        tb200_plan_group &G = R.groups[gm.first];
        const struct ggml_tensor *other_t = mm == n->src_[0] ? (const struct ggml_tensor *)n->src_[1] : (const struct ggml_tensor *)n->src_[0];
        const bool same_group = slot.count(other_t) && slot[other_t].first == gm.first;   // a sum of two matmuls of ONE group: no barrier stands between them
        if (n_uses(mm) != 1 || is_output(tv(mm)) || G.residual[gm.second] || other->ne[1] != 1 || tv(mm)->op_ != TB200_GGML_OP_MUL_MAT || same_group) break;
        G.out[gm.second] = (float *)n->data, G.residual[gm.second] = (const float *)other->data;
        R.folded++;
    }
    // every folded intermediate must be read by its folding nodes only
    for (auto &kv : virt) {
        const Virt &v = kv.second;
        if (v.fused_uses != n_uses(kv.first) || v.fused_uses == 0 || is_output(tv((const struct ggml_tensor *)kv.first))) {
            if (v.start < cut) cut = v.start;
        } else R.folded++;
    }
    if (cut < limit) return false;
    R.end = j;
    return true;
}

int backend_graph_compute(tb200_backend_t bk, struct ggml_cgraph *cg_) {
    BackendCtx *c = (BackendCtx *)bk->context;
    const tb200_ggml_cgraph_view *cg = (const tb200_ggml_cgraph_view *)cg_;
    DeviceGuard g(c->device);
    tb200_ctx *mc = tb200_get_ctx();
    if (!mc) return TB200_GGML_STATUS_FAILED;
    struct StreamSwap {   // the module's device-level calls run on the context's current stream: point it at this backend's for the graph
        tb200_ctx *mc; cudaStream_t old;
        StreamSwap(tb200_ctx *m, cudaStream_t s) : mc(m), old(m->stream) { m->stream = s; }
        ~StreamSwap() { mc->stream = old; }
    } swap(mc, c->stream);
    const uint64_t l0 = tb200_launch_count();
    c->last_plan_nodes = 0;
    int rc = TB200_OK;
    bool has_ew = false;
    for (int q = 0; q < cg->n_nodes && !has_ew; ++q) has_ew = is_ew_op(tv(cg->nodes[q])->op_);
    UseMap uses;   // filled on the first plan that has to be built for a graph with elementwise nodes
    for (int i = 0; i < cg->n_nodes && rc == TB200_OK;) {
        const T *op = tv(cg->nodes[i]);
        if (is_view_op(op->op_)) { ++i; continue; }
        // a run of consecutive one-token Q4_0 (or Q4_K) matmuls and their foldable neighbours: ONE persistent kernel for the whole run
        // (matmuls sharing their input form a group)
        const bool starts_run = plan_node(op) != 0 || (has_ew && is_ew_op(op->op_) && ew_supported(op) && op->ne[1] == 1 &&
                                                      (op->op_ == TB200_GGML_OP_RMS_NORM || op->op_ == TB200_GGML_OP_UNARY));
        if (starts_run) {
            // cheap extent of what COULD be one run (no folding rules yet): the plan cache is keyed by these nodes' pointers
            int jmax = i;
            PlanKey key;
            const int type0 = plan_node(op);
            for (; jmax < cg->n_nodes; ++jmax) {
                const T *n = tv(cg->nodes[jmax]);
                if (is_view_op(n->op_)) continue;
                const int t = plan_node(n);
                if (!(t != 0 || (has_ew && is_ew_op(n->op_) && ew_supported(n) && n->ne[1] == 1))) break;
                if (t && type0 && t != type0) break;
                key.ptrs.push_back(n->data);
                key.ptrs.push_back(n->src_[0] ? tv((const struct ggml_tensor *)n->src_[0])->data : nullptr);
                key.ptrs.push_back(n->src_[1] ? tv((const struct ggml_tensor *)n->src_[1])->data : nullptr);
                // op, its first parameter word (rms_norm's eps, the unary function) and the row length: an allocator may hand the same
                // addresses to a different graph
                key.ptrs.push_back((const void *)((uintptr_t)(uint32_t)n->op_ | ((uintptr_t)(uint32_t)n->op_params_[0] << 32)));
                key.ptrs.push_back((const void *)(uintptr_t)n->ne[0]);
            }
            auto it = c->plans.find(key);
            if (it != c->plans.end() && it->second.epoch != g_resident_epoch.load()) {   // some resident weights were freed since: the plan may hold one
                cudaStreamSynchronize(c->stream);
                if (it->second.plan) tb200_plan_free(it->second.plan);
                c->plans.erase(it);
                it = c->plans.end();
            }
            if (it == c->plans.end()) {
                if (has_ew && uses.empty())
                    for (int q = 0; q < cg->n_nodes; ++q)
                        for (int sidx = 0; sidx < 10; ++sidx)
                            if (tv(cg->nodes[q])->src_[sidx]) uses[tv(cg->nodes[q])->src_[sidx]].push_back(q);
                FusedRun R;
                int limit = jmax;
                bool ok = false;
                while (limit > i) {
                    int cut = limit;
                    if (build_fused_run(cg, i, limit, uses, R, cut)) { ok = true; break; }
                    limit = cut < limit ? cut : limit - 1;
                }
                ok = ok && R.end > i && (R.matmuls >= 2 || (R.matmuls >= 1 && R.folded > 0));
                tb200_plan *p = ok ? tb200_plan_create((int)R.groups.size(), R.groups.data()) : nullptr;
                if (c->plans.size() >= 512) {   // graphs that keep changing: do not hoard plans (a 80-layer model is ~160 runs per graph)
                    cudaStreamSynchronize(c->stream);
                    for (auto &kv : c->plans) if (kv.second.plan) tb200_plan_free(kv.second.plan);
                    c->plans.clear();
                }
                it = c->plans.emplace(key, BackendCtx::PlanEntry{p, p ? (int)R.groups.size() : 0, g_resident_epoch.load(), p ? R.end : i, p ? R.matmuls : 0}).first;   // plan == nullptr: "cannot be planned", node by node below
            }
            if (it->second.plan) {
                rc = tb200_plan_run(it->second.plan, 0, it->second.groups);
                i = it->second.end;
                c->last_plan_nodes += it->second.matmuls;
                continue;
            }
        }
        if (op->op_ == TB200_GGML_OP_MUL_MAT && mul_mat_supported(op)) rc = run_mul_mat(c, op);
        else if (op->op_ == TB200_GGML_OP_MUL_MAT_ID && mul_mat_id_supported(op)) rc = run_mul_mat_id(c, op);
        else if (is_ew_op(op->op_) && ew_supported(op)) rc = run_ew(op);
        else {
            char msg[128];
            snprintf(msg, sizeof msg, "tb200_backend: node %d (op %d) is not supported by this backend\n", i, op->op_);
            backend_log(msg);
            return TB200_GGML_STATUS_FAILED;
        }
        ++i;
    }
    c->last_launches = (int)(tb200_launch_count() - l0);
    if (rc != TB200_OK) {
        char msg[600];
        snprintf(msg, sizeof msg, "tb200_backend: graph_compute failed: %s\n", tb200_last_error());
        backend_log(msg);
        return TB200_GGML_STATUS_FAILED;
    }
    return TB200_GGML_STATUS_SUCCESS;
}

const tb200_ggml_backend_i kBackendIface = {
    backend_get_name, backend_free, backend_default_buft, backend_set_tensor_async, backend_get_tensor_async,
    /* cpy_tensor_async */ nullptr, backend_synchronize,
    /* graph_plan_create / free / update / compute */ nullptr, nullptr, nullptr, nullptr,
    backend_graph_compute, backend_supports_op, backend_supports_buft, backend_offload_op,
    /* events */ nullptr, nullptr, nullptr, nullptr, nullptr,
};

tb200_backend_t reg_init(const char *, void *user_data) { return tb200_backend_init((int)(intptr_t)user_data); }

}  // namespace

extern "C" {

// Stand-alone host (api == NULL: no ggml in the process, e.g. bench.py driving the backend object directly): the two host
// services the module needs, in their reference form (ggml_backend_buffer_init: ggml-backend.c:60-80; ggml_nbytes: ggml.c:3500-3520).
static tb200_buffer_t standalone_buffer_init(tb200_buft_t buft, struct tb200_ggml_backend_buffer_i iface, void *context, size_t size) {
    tb200_ggml_backend_buffer *b = (tb200_ggml_backend_buffer *)malloc(sizeof(*b));
    b->iface = iface, b->buft = buft, b->context = context, b->size = size, b->usage = TB200_GGML_BUFFER_USAGE_ANY;
    return b;
}
static size_t standalone_nbytes(const struct ggml_tensor *t_) {
    const T *t = tv(t_);
    const long blck = t->type == TB200_TYPE_F32 || t->type == TB200_TYPE_F16 || t->type == TB200_TYPE_I32 ? 1 : (tb200_is_kquant(t->type) || t->type == TB200_TYPE_Q8_K ? QK_K : QK);
    size_t n = (size_t)(t->ne[0] / blck) * t->nb[0];
    for (int i = 1; i < 4; ++i) n += (size_t)(t->ne[i] - 1) * t->nb[i];
    return n;
}

bool tb200_backend_link(const struct tb200_ggml_backend_api *api) {
    if (api && (!api->ggml_backend_buffer_init || !api->ggml_nbytes)) { tb200_set_error("backend_link: the host API table is missing required entries"); return false; }
    if (api) g_api = *api;
    else {
        memset(&g_api, 0, sizeof g_api);
        g_api.malloc = malloc, g_api.free = free, g_api.ggml_backend_buffer_init = standalone_buffer_init, g_api.ggml_nbytes = standalone_nbytes;
    }
    tb200_host_api h = {g_api.malloc, g_api.free, nullptr, nullptr};
    if (!tb200_link(&h)) return false;
    const int n = tb200_device_count();
    for (int d = 0; d < n && d < kMaxDev; ++d) {
        g_buft_ctx[d].device = d;
        snprintf(g_buft_ctx[d].name, sizeof g_buft_ctx[d].name, "TB200_%d", d);
        g_buft[d].iface = {buft_get_name, buft_alloc_buffer, buft_get_alignment, /* get_max_size */ nullptr, /* get_alloc_size */ nullptr, buft_is_host};
        g_buft[d].context = &g_buft_ctx[d];
    }
    g_host_buft.iface = {host_buft_get_name, host_buft_alloc_buffer, host_buft_get_alignment, nullptr, nullptr, host_buft_is_host};
    g_host_buft.context = nullptr;
    g_linked = true;
    return n > 0;
}

tb200_backend_t tb200_backend_init(int device) {
    if (!g_linked || device < 0 || device >= tb200_device_count()) { tb200_set_error("backend_init: invalid device %d (or tb200_backend_link not called)", device); return nullptr; }
    DeviceGuard g(device);
    BackendCtx *c = new BackendCtx();
    c->device = device;
    snprintf(c->name, sizeof c->name, "TB200_%d", device);
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) { cudaGetLastError(); delete c; return nullptr; }
    tb200_ggml_backend *b = new tb200_ggml_backend();
    b->guid = g_guid, b->iface = kBackendIface, b->context = c;
    return b;
}
bool tb200_backend_is_tb200(tb200_backend_t b) { return b && b->guid && !memcmp(b->guid, g_guid, 16); }
tb200_buft_t tb200_backend_buffer_type(int device) { return g_linked && device >= 0 && device < tb200_device_count() && device < kMaxDev ? &g_buft[device] : nullptr; }
tb200_buft_t tb200_backend_host_buffer_type(void) { return g_linked && g_api.ggml_backend_cpu_buffer_from_ptr ? &g_host_buft : nullptr; }
// ggml's split buffer type shards a weight matrix by rows over the devices of ONE process (ggml-cuda.cu:646-990). This module
// runs one process per GPU and shards through its peer API (tb200_mul_mat_acts_multi_gather), so there is no such buffer type.
tb200_buft_t tb200_backend_split_buffer_type(const float *) { return nullptr; }
int tb200_backend_reg_devices(void) {
    if (!g_linked || !g_api.ggml_backend_register) return 0;
    const int n = tb200_device_count();
    for (int d = 0; d < n; ++d) g_api.ggml_backend_register(g_buft_ctx[d].name, reg_init, &g_buft[d], (void *)(intptr_t)d);
    return n;
}
int tb200_backend_get_device_count(void) { return tb200_device_count(); }
void tb200_backend_get_device_memory(int device, size_t *free_bytes, size_t *total_bytes) {
    DeviceGuard g(device);
    if (cudaMemGetInfo(free_bytes, total_bytes) != cudaSuccess) { cudaGetLastError(); *free_bytes = *total_bytes = 0; }
}
void tb200_backend_get_device_description(int device, char *description, size_t description_size) {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, device) != cudaSuccess) { cudaGetLastError(); snprintf(description, description_size, "unknown"); return; }
    snprintf(description, description_size, "%s", p.name);
}
void tb200_backend_get_device_properties(int device, struct tb200_device_properties *out) {
    cudaDeviceProp p;
    memset(out, 0, sizeof *out);
    if (cudaGetDeviceProperties(&p, device) != cudaSuccess) { cudaGetLastError(); return; }
    snprintf(out->name, sizeof out->name, "%s", p.name);
    out->totalGlobalMem = p.totalGlobalMem, out->multiProcessorCount = p.multiProcessorCount, out->major = p.major, out->minor = p.minor;
    snprintf(out->compute, sizeof out->compute, "%d%d", p.major, p.minor);
}
bool tb200_backend_register_host_buffer(void *buffer, size_t size) {
    if (cudaHostRegister(buffer, size, cudaHostRegisterPortable | cudaHostRegisterReadOnly) != cudaSuccess) { cudaGetLastError(); return false; }
    return true;
}
void tb200_backend_unregister_host_buffer(void *buffer) { if (cudaHostUnregister(buffer) != cudaSuccess) cudaGetLastError(); }
void tb200_backend_last_graph_stats(tb200_backend_t b, int *launches, int *plan_nodes) {
    BackendCtx *c = (BackendCtx *)b->context;
    if (launches) *launches = c->last_launches;
    if (plan_nodes) *plan_nodes = c->last_plan_nodes;
}

}  // extern "C"
