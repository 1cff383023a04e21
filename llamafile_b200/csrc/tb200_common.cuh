// tinyBLAS-B200 internal definitions shared by all translation units (sm_100a only).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/tinyblas_b200.h"

#ifndef __CUDA_ARCH__
#define TB200_HOST_ONLY 1
#endif
#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "tinyBLAS-B200 is written for sm_100a only"
#endif

#define QK 32
#define QK_K 256

static_assert(sizeof(tb200_block_q4_0) == 18, "block_q4_0");
static_assert(sizeof(tb200_block_q8_0) == 34, "block_q8_0");
static_assert(sizeof(tb200_block_q4_K) == 144, "block_q4_K");
static_assert(sizeof(tb200_block_q8_K) == 292, "block_q8_K");
static_assert(sizeof(tb200_block_q6_K) == 210, "block_q6_K");
static_assert(sizeof(tb200_block_q4_1) == 20 && sizeof(tb200_block_q5_0) == 22 && sizeof(tb200_block_q8_1) == 36, "block_q4_1 / q5_0 / q8_1");
static_assert(sizeof(tb200_block_q5_K) == 176 && sizeof(tb200_block_iq4_xs) == 136, "block_q5_K / iq4_xs");

// weight types whose activations are Q8_K (256-element super-blocks)
static inline bool tb200_is_kquant(int type) { return type == TB200_TYPE_Q4_K || type == TB200_TYPE_Q6_K || type == TB200_TYPE_Q5_K || type == TB200_TYPE_IQ4_XS; }
// weight types held in the generic resident form G8 (below): the remaining types iqk serves (llamafile/iqk_mul_mat.inc:1408-1463)
static inline bool tb200_is_g8(int type) { return type == TB200_TYPE_Q4_1 || type == TB200_TYPE_Q5_0 || type == TB200_TYPE_Q5_K || type == TB200_TYPE_IQ4_XS; }
#define TB200_TYPE_G8 1000   // kernel-template tag of that form (never a ggml type code)

// ---------------------------------------------------------------------------------------------
// Resident weight matrix: the GGUF rows split into 16-byte-aligned planes (one-time repack at
// upload, tb200_repack.cu). Total bytes are identical to the GGUF encoding.
//   Q4_0: qs  [rows][k/32] x 16 B  (byte j: low nibble = elem j, high nibble = elem j+16)
//         sc  [rows][k/32] fp16 d
//   Q8_0: qs  [rows][k] int8 ;  sc [rows][k/32] fp16 d
//   Q4_K: qs  [rows][k/256] x 128 B ; sc [rows][k/256] x 16 B = {fp16 d, fp16 dmin, u8 scales[12]}
//   Q6_K: qs  [rows][k/256] x 192 B = 4 units x 3 pieces of 16 B (unit u = 2*half + l/16 covers elements
//              128*half + 16*(l/16) + 32*j + [0,16), j = 0..3: one 64-wide K stage of the prefill path, one lane of the decode
//              path); piece j < 3, byte e = q_j[e] << 2 | two bits of q_3[e] (bits 5:4, 3:2, 1:0 for j = 0, 1, 2)
//         sc  [rows][k/256] x 16 B int8 scales, byte 4u + j ; d [rows][k/256] fp16
//   G8  : (Q4_1, Q5_0, Q5_K, IQ4_XS) every 32-block as int8 quants q and an fp16 pair {a, b}: value = a*q + b, exactly what the
//         block decodes to - Q4_1: q = nibble, a = d, b = m; Q5_0: q = q5 - 16, a = d; Q5_K: q = q5, a = fp16(d*sc), b = fp16(-dmin*m);
//         IQ4_XS: q = kvalues_iq4nl[nibble], a = fp16(d*(ls - 32)). qs [rows][k] int8 ; sc [rows][k/32] x {fp16 a, fp16 b}.
//         1.125 B/element in HBM instead of 0.53..0.69: these types are served (decode and prefill, bit-for-bit integer block
//         dots), at about half the bandwidth efficiency of a native nibble kernel.
//   F16 : qs  [rows][k] fp16 ; sc = nullptr
// ---------------------------------------------------------------------------------------------
struct tb200_weights {
    int type;
    long rows, k;
    void *qs;
    void *sc;
    void *d;            // Q6_K only: per-super-block fp16 scale plane
    size_t qs_row_bytes, sc_row_bytes, d_row_bytes;
    size_t bytes;
    int device;
    void *tmap_cache;   // prefill GEMM tensor maps (tb200_gemm.cu), lazily built
};

// Quantized activations in the module's internal planar form (one per call, lives in scratch):
//   q   int8  [tokens][k]
//   ds  {float d, int32 sum} per 32-block [tokens][k/32]   (Q8_0: d = fp16 scale widened; sum = sum of q)
//   Q8_K: d is the per-256 float scale replicated on its 8 sub-blocks, sum = bsums[2s]+bsums[2s+1]

// peer group for the fused all-gather (tb200_peer.cu)
struct tb200_peers {
    int world, rank;
    void *out_base[TB200_MAX_PEERS];    // symmetric output buffer on every rank (own entry = local pointer)
    uint32_t *flags[TB200_MAX_PEERS];   // [world] arrival epochs on every rank
    uint32_t *epoch;                    // local launch counter (device memory, so graph replays advance it)
    int *timeout_flag;                  // set when a barrier gave up waiting
    uint32_t *plan_seq;                 // shared groups completed by ALL multi-GPU plan launches on this peer group (device memory; the plans' flag values)
};

// per-device context (tb200_api.cu)
struct tb200_ctx {
    int device;
    int sm_count;
    cudaStream_t stream;
    void *scratch;          // device scratch arena (activations, gathered rows, ...)
    size_t scratch_bytes;
    void *pinned;           // host staging for the host-pointer entry points
    size_t pinned_bytes;
    void *scratch2;         // second arena: fp32 intermediates of the fused neighbours (tb200_fused.cu), alive across a matmul that uses `scratch`
    size_t scratch2_bytes;
};

tb200_ctx *tb200_get_ctx();                       // nullptr when no device
void *tb200_scratch(tb200_ctx *c, size_t bytes);  // grows on demand (stream-ordered)
void *tb200_scratch2(tb200_ctx *c, size_t bytes);
void tb200_set_error(const char *fmt, ...);
void tb200_count_launch(int n = 1);

#define TB_CUDA(expr)                                                                              \
    do {                                                                                           \
        cudaError_t e_ = (expr);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            tb200_set_error("%s:%d %s: %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e_));    \
            return TB200_ERR_CUDA;                                                                 \
        }                                                                                          \
    } while (0)

// ---- kernels' launchers (each defined in its own .cu) ------------------------------------------
// quantizers: raw GGUF blocks and/or planar activations
int tb200_launch_quantize_q8_0(tb200_ctx *c, const float *x, long ldx, long rows, long k, void *raw, int8_t *q, int2 *ds);
int tb200_launch_quantize_q8_K(tb200_ctx *c, const float *x, long ldx, long rows, long k, void *raw, int8_t *q, int2 *ds);
int tb200_launch_f32_to_f16(tb200_ctx *c, const float *x, long ldx, long rows, long k, __half *y);
// raw GGUF activation blocks -> planar
int tb200_launch_unpack_q8_0(tb200_ctx *c, const void *raw, long ldb_blocks, long rows, long k, int8_t *q, int2 *ds);
int tb200_launch_unpack_q8_K(tb200_ctx *c, const void *raw, long ldb_blocks, long rows, long k, int8_t *q, int2 *ds);
int tb200_launch_unpack_q8_1(tb200_ctx *c, const void *raw, long ldb_blocks, long rows, long k, int8_t *q, int2 *ds);
// repack
int tb200_launch_repack(tb200_ctx *c, const void *raw, int type, long rows, long k, size_t row_stride, void *qs, void *sc, void *d);
// decode path. out[t*ldo + i]. rowmap: optional per-token (weight-row offset, output pointer) for mixmul.
// grouped decode (mixmul): one item = rows [row0, row0+nrows) of one expert against ntok gathered tokens
struct tb200_gemv_item {
    long row0;
    int ntok;
    int a_row[TB200_GEMV_MAX_TOKENS];     // activation row index per gathered token
    int out_row[TB200_GEMV_MAX_TOKENS];   // output row index (out + out_row*ldo)
};
struct tb200_gemv_args {
    const tb200_weights *w;
    long row0, nrows;              // weight rows [row0, row0+nrows)
    const int8_t *aq; const int2 *ads;    // planar activations (quantized types)
    const void *af;                // F16 weights: activations as fp16 (tokens>2) or f32 (tokens<=2)
    int af_is_f32;
    long a_stride;                 // activation row stride in elements
    long tokens;
    float *out; long ldo;
    const tb200_gemv_item *items; long n_items;   // optional grouped mode (device array); tokens = template T
    const int *n_live;             // optional with items: device count of live items, which must lead the table (resident grouped kernel)
    // optional (Q4_0 / Q4_K): more matrices sharing K and the activations, served by the same launch
    int nmat; const tb200_weights *w_more[TB200_GEMV_MAX_MATS - 1]; float *out_more[TB200_GEMV_MAX_MATS - 1]; long ldo_more[TB200_GEMV_MAX_MATS - 1];
    // optional (Q4_0 / Q4_K): fused all-gather — `out` pointers lie inside this rank's symmetric buffer `local_base`, and
    // every result is stored at the same offset of every rank's copy (peer_base[0..world))
    int world; const void *local_base; void *peer_base[TB200_MAX_PEERS];
};
int tb200_launch_gemv(tb200_ctx *c, const tb200_gemv_args &a);
int tb200_launch_blockdot(tb200_ctx *c, const tb200_weights *w, const int8_t *aq, const int2 *ads, int32_t *out);

bool tb200_pdl_enabled();   // TB200_PDL=0 disables programmatic dependent launch

#ifdef __CUDACC__
// Every kernel of the module is launched with programmatic stream serialization (PDL): it may start while its
// predecessor in the stream is still draining. Contract inside the kernels: pdl_launch_dependents() first (lets
// the successor be scheduled early), immutable data (resident weights) may be read at once, and pdl_wait() must
// precede the first access to anything a predecessor may have written or may still be reading.
template <typename... KArgs, typename... Args>
inline cudaError_t tb200_launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid, cfg.blockDim = block, cfg.dynamicSmemBytes = smem, cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr, cfg.numAttrs = tb200_pdl_enabled() ? 1 : 0;
    tb200_count_launch();
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// ---- device helpers ----------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ uint4 ldg_stream(const uint4 *p) {
    uint4 r;   // weights are read exactly once: bypass L1 allocation
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ int dp4a_ss(int a, int b, int c) { return __dp4a(a, b, c); }
__device__ __forceinline__ int dp4a_us(unsigned a, int b, int c) {   // unsigned bytes x signed bytes
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
#endif
