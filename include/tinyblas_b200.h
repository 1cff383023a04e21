/* tinyBLAS-B200 — C ABI of the B200-native quantized-matmul module.
 *
 * The module is one shared object, libtinyblas_b200.so (sm_100a CUDA + a C++ host layer), loaded
 * the way llamafile loads its GPU DSO (llamafile/cuda.c:701-753: dlopen -> dlsym table -> link()
 * handshake that answers "is there a usable device"). Nothing in these signatures is a torch or
 * C++ type: plain pointers, sizes and ggml type codes only.
 *
 * Part 1 are the reference's own entry points for the hot path (llamafile/sgemm.h:23-28, 76-77),
 * same names, argument meaning and error behaviour, taking HOST pointers exactly like the CPU
 * implementation they replace. Part 2 is the additive device-level API a GPU backend needs
 * (last row): resident weights, device-pointer matmuls, activation quantizers.
 *
 * There is no CPU fallback anywhere in the module: when no sm_100 device is present
 * tb200_link() returns false and every compute entry point returns false / TB200_ERR_NO_DEVICE.
 */
#ifndef TINYBLAS_B200_H
#define TINYBLAS_B200_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* llamafile is a Cosmopolitan binary and marks every cross-module pointer ms_abi
 * (llama.cpp/ggml/include/ggml.h:190-198). Our host is glibc; keep the hook. */
#ifndef TB200_CALL
#define TB200_CALL
#endif
#define TB200_API __attribute__((visibility("default")))

/* ggml type codes (llama.cpp/ggml/include/ggml.h:361-397) */
enum tb200_type {
    TB200_TYPE_F32 = 0,
    TB200_TYPE_F16 = 1,
    TB200_TYPE_Q4_0 = 2,
    TB200_TYPE_Q4_1 = 3,    /* the remaining types iqk serves on the same entry points */
    TB200_TYPE_Q5_0 = 6,
    TB200_TYPE_Q8_0 = 8,
    TB200_TYPE_Q8_1 = 9,    /* activations of Q4_1 (vec_dot_type) */
    TB200_TYPE_Q4_K = 12,
    TB200_TYPE_Q5_K = 13,
    TB200_TYPE_Q6_K = 14,   /* the type Q4_K_M files keep output / ffn_down / attn_v in */
    TB200_TYPE_Q8_K = 15,
    TB200_TYPE_IQ4_XS = 23,
    TB200_TYPE_I32 = 26,
};

/* GGUF block layouts, unchanged (llama.cpp/ggml/src/ggml-common.h:144-149, 186-191, 266-277;
 * block_q8_K in llamafile's field order {d, bsums, qs}: llama.cpp.patches/patches/ggml-common.h.patch:25-35) */
#pragma pack(push, 1)
typedef struct { uint16_t d; uint8_t qs[16]; } tb200_block_q4_0;
typedef struct { uint16_t d; int8_t qs[32]; } tb200_block_q8_0;
typedef struct { uint16_t d, dmin; uint8_t scales[12]; uint8_t qs[128]; } tb200_block_q4_K;
typedef struct { float d; int16_t bsums[16]; int8_t qs[256]; } tb200_block_q8_K;
typedef struct { uint8_t ql[128]; uint8_t qh[64]; int8_t scales[16]; uint16_t d; } tb200_block_q6_K;   /* ggml-common.h:299-307 */
typedef struct { uint16_t d, m; uint8_t qs[16]; } tb200_block_q4_1;                                     /* ggml-common.h:151-161 */
typedef struct { uint16_t d; uint8_t qh[4]; uint8_t qs[16]; } tb200_block_q5_0;                          /* ggml-common.h:164-169 */
typedef struct { uint16_t d, s; int8_t qs[32]; } tb200_block_q8_1;                                       /* ggml-common.h:193-203 */
typedef struct { uint16_t d, dmin; uint8_t scales[12]; uint8_t qh[32]; uint8_t qs[128]; } tb200_block_q5_K;   /* ggml-common.h:280-295 */
typedef struct { uint16_t d; uint16_t scales_h; uint8_t scales_l[4]; uint8_t qs[128]; } tb200_block_iq4_xs;   /* ggml-common.h:393-398 */
#pragma pack(pop)

/* Layout-compatible prefix views of the two ggml structs llamafile_mixmul reads
 * (struct ggml_tensor: llama.cpp/ggml/include/ggml.h:583-618, fields type/ne/nb/data;
 *  struct ggml_compute_params: llama.cpp.patches/patches/ggml.h.patch:48-57). */
struct ggml_tensor;
struct ggml_compute_params;
typedef struct {
    int32_t type;
    int32_t backend_;
    void *buffer_;
    int64_t ne[4];
    size_t nb[4];
    int32_t op_;
    int32_t op_params_[16];
    int32_t flags_;
    void *grad_;
    void *src_[10];
    void *view_src_;
    size_t view_offs_;
    void *data;
    /* char name[128]; void *extra;  (not read) */
} tb200_ggml_tensor_view;
typedef struct {
    int ith, nth;
    size_t wsize;
    void *wdata;
    void *shared_;
} tb200_ggml_compute_params_view;

/* ------------------------------------------------------------------------------------------
 * Part 1 — the reference's entry points (host pointers)
 * ------------------------------------------------------------------------------------------ */

/* llamafile/sgemm.h:23-24, llamafile/sgemm.cpp:128-131. C[ldc*j + i] = sum_l A[lda*i + l] * B[ldb*j + l],
 * m = weight rows, n = tokens, k/lda/ldb counted in BLOCKS of Atype/Btype, ldc in floats.
 * Serviced (returns true, C written) for the combinations the reference's x86 build services:
 *   Q4_0 x Q8_0, Q8_0 x Q8_0, Q4_K x Q8_K, Q6_K x Q8_K, Q5_K x Q8_K, IQ4_XS x Q8_K, Q5_0 x Q8_0, Q4_1 x Q8_1,
 *   F16 x F16, F16 x F32 (n <= 2); Ctype F32. (F32 and BF16
 *   weights, which the reference also serves, are outside north_star's list and are handed back.)
 * Anything else returns false before touching C ("not handled, caller falls back"), never an error.
 * Threading contract: ggml calls this from all nth worker threads. Thread 0 does the whole job
 * on the GPU and returns after C is complete; threads ith != 0 return true immediately and rely on
 * the caller's barrier after the op (ggml.c.patch:2316), as the CPU kernels do for their own tiles. Because the other
 * threads have answered before thread 0 touches the GPU, the true/false answer is a pure function of the arguments plus a
 * sticky "module disabled" flag: a runtime failure on thread 0 is fatal for the call in flight (logged, abort() unless
 * TB200_ABORT_ON_ERROR=0 or nth == 1) and every later call returns false on every thread.
 * A is uploaded once and kept ONLY when it lies inside a range registered with tb200_register_weights() (the mmap'd
 * GGUF file, immutable); any other A (KV cache, scratch) is uploaded on every call and dropped after it.
 * One host-pointer call runs at a time per device (internal lock); callers on different devices run concurrently. */
TB200_API bool llamafile_sgemm(long m, long n, long k, const void *A, long lda, const void *B, long ldb,
                               void *C, long ldc, int ith, int nth, int Atype, int Btype, int Ctype);

/* llamafile/sgemm.h:25-28, llamafile/tinyblas_cpu_mixmul.inc:390-398. weights [cols, rows, experts],
 * thought [cols, tasks, tokens] (F32 or the vec_dot type), plan [thinkers, tokens] I32,
 * result [rows, thinkers, tokens] F32. Weight types F16, Q8_0, Q4_0 as in the reference; false otherwise. */
TB200_API bool llamafile_mixmul(const struct ggml_compute_params *params, const struct ggml_tensor *weights,
                                const struct ggml_tensor *thought, const struct ggml_tensor *plan,
                                struct ggml_tensor *result);

/* llamafile/tinyblas_cpu_mixmul_amd_avx.cpp:8-18: host scratch bytes llamafile_mixmul would carve from
 * params->wdata. The GPU module keeps its scratch in HBM, but reports the reference's figure so that
 * ggml_graph_plan (ggml.c.patch:2279-2285) sizes wdata identically. */
TB200_API size_t llamafile_mixmul_needs(const struct ggml_tensor *weights, const struct ggml_tensor *thought,
                                        const struct ggml_tensor *plan);

/* llamafile/sgemm.h:76-77, llamafile/iqk_mul_mat.inc:204-221: one expert's rows with an
 * mmid_row_mapping {int32 i1, int32 i2}[Ny]; typeA one of the k-quants held here (Q4_K, Q5_K, Q6_K, IQ4_XS; B rows Q8_K), Q4_0 or Q5_0
 * (B rows Q8_0) or Q4_1 (B rows Q8_1); false for anything else. */
TB200_API bool llamafile_mixmul_iqk(long Nx, long Ny, long ne00, int ne11, int typeA, const void *A,
                                    const void *B, float *C, long nb1, long nb2, const void *vrow_mapping,
                                    int ith, int nth);

/* Declares host range [base, base+bytes) immutable until tb200_unregister_weights(base): weight matrices inside it are
 * uploaded (and repacked) once, keyed by (A, m, k, lda, Atype), and stay resident. llamafile would call this right after
 * mmap()ing the GGUF file (llama.cpp/src/llama.cpp llama_mmap) and unregister before munmap. Unregistering frees the
 * resident copies made from the range. */
TB200_API int tb200_register_weights(const void *base, size_t bytes);
TB200_API int tb200_unregister_weights(const void *base);
TB200_API void tb200_weights_cache_clear(void);      /* drops every resident copy (ranges stay registered) */
TB200_API size_t tb200_weights_cache_bytes(void);    /* HBM bytes held by resident copies */

/* from_float of the vec_dot types (ggml.c type_traits: quantize_row_q8_0 ggml-quants.c:866/940-990,
 * quantize_row_q8_K :3593-3645, ggml_fp32_to_fp16_row). Host in, host out, bit-exact. */
TB200_API bool tb200_quantize_row_q8_0(const float *x, void *y, int64_t k);
TB200_API bool tb200_quantize_row_q8_K(const float *x, void *y, int64_t k);
TB200_API bool tb200_fp32_to_fp16_row(const float *x, uint16_t *y, int64_t k);

/* ------------------------------------------------------------------------------------------
 * Part 2 — module handshake and device-level API (additive)
 * ------------------------------------------------------------------------------------------ */

enum tb200_status {
    TB200_OK = 0,
    TB200_ERR_NO_DEVICE = 1,
    TB200_ERR_UNSUPPORTED = 2,   /* the reference would return false for this request */
    TB200_ERR_INVALID = 3,
    TB200_ERR_CUDA = 4,
};

/* Host services handed to the module at link time, the way llamafile passes libc + ggml helpers
 * through struct ggml_backend_api (llama.cpp.patches/patches/ggml-backend-impl.h.patch:20-58).
 * Every member may be NULL (the module then uses its own libc). */
struct tb200_host_api {
    void *(TB200_CALL *malloc)(size_t);
    void (TB200_CALL *free)(void *);
    long (TB200_CALL *write)(int fd, const void *buf, size_t n);   /* log sink, as g_backend->write */
    const char *(TB200_CALL *getenv)(const char *);
};

/* Handshake, mirrors ggml_cuda_link (ggml-cuda.cu.patch:468-520): stores the table, initialises the
 * current device, returns true iff an sm_100 device is usable. */
TB200_API bool tb200_link(const struct tb200_host_api *api);
TB200_API int tb200_device_count(void);
TB200_API int tb200_set_device(int device);
TB200_API const char *tb200_last_error(void);
TB200_API int tb200_sm_count(void);
/* Stream used by all device-level calls of the calling thread's device (cudaStream_t as void*). */
TB200_API int tb200_set_stream(void *cuda_stream);
TB200_API int tb200_synchronize(void);

/* Resident weights. `rows` weight rows of `k` ELEMENTS each in GGUF layout (F16, Q8_0, Q4_0, Q4_K),
 * `row_stride` bytes apart. Upload = copy + one-time device repack into aligned planes (quants plane,
 * scales plane) of identical total size; the GGUF bytes themselves are not modified. */
typedef struct tb200_weights tb200_weights;
TB200_API tb200_weights *tb200_weights_upload(const void *host_rows, int type, long rows, long k, size_t row_stride);
TB200_API tb200_weights *tb200_weights_from_device(const void *dev_rows, int type, long rows, long k, size_t row_stride);
TB200_API void tb200_weights_free(tb200_weights *w);
TB200_API size_t tb200_weights_bytes(const tb200_weights *w);

/* Device-pointer activation quantizers: x [rows][k] f32 (row stride ldx floats) -> GGUF blocks,
 * rows contiguous. Bit-exact to the reference. */
TB200_API int tb200_quantize_q8_0_dev(const float *x, long ldx, void *y, long rows, long k);
TB200_API int tb200_quantize_q8_K_dev(const float *x, long ldx, void *y, long rows, long k);

/* ggml_compute_forward_mul_mat for one weight matrix (ggml.c:12277-12403): out[t*ldo + i] =
 * W[i] . vec_dot_type(x[t]); x f32 [tokens][k] on the device, quantized on the device exactly as the
 * reference does (Q8_0 / Q8_K / F16, or kept F32 for F16 weights when tokens <= 2). Decode kernels for
 * tokens <= TB200_GEMV_MAX_TOKENS, tcgen05 GEMM above. */
#define TB200_GEMV_MAX_TOKENS 8
#define TB200_GEMV_MAX_MATS 4
#define TB200_MAX_PEERS 8
TB200_API int tb200_mul_mat(const tb200_weights *w, const float *x, long ldx, long tokens, float *out, long ldo);

/* tb200_mul_mat for callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc). NO COUNTERPART IN THE REFERENCE. Same arguments, same kernels and
 * bit-identical results as tb200_mul_mat, but the quantized activations go into a fixed device array inside the library
 * (TB200_NOALLOC_SCRATCH_BYTES of HBM per GPU that loads the library, held until exit) instead of the context's growable
 * cudaMalloc arena, and the call path contains no cudaMalloc / cudaFree / allocating library routine.
 * Limits: tokens <= TB200_GEMV_MAX_TOKENS (decode-size calls) and tokens * k * 1.25 bytes (quantized weight types; tokens * k * 2
 * for F16 weights with 3..8 tokens), rounded up to 256, <= TB200_NOALLOC_SCRATCH_BYTES - e.g. 8 tokens at k = 32768; anything
 * larger returns TB200_ERR_INVALID with tb200_last_error() naming the limit, before any launch.
 * The array is shared by every call on the GPU and keeps the last call's contents (nothing in it needs to start at zero: the
 * quantizer rewrites all the decode kernel reads): calls that OVERLAP on the GPU - two streams, two graph branches - race on
 * it; calls in one stream are ordered. The first call per device resolves the array's address (cudaGetSymbolAddress, no
 * allocation); make that first call outside a stream capture. */
#define TB200_NOALLOC_SCRATCH_BYTES (320 * 1024)
TB200_API int tb200_mul_mat_noalloc(const tb200_weights *w, const float *x, long ldx, long tokens, float *out, long ldo);

/* Decode only (tokens <= TB200_GEMV_MAX_TOKENS): n <= TB200_GEMV_MAX_MATS weight matrices of the same type and K that
 * consume the SAME activations (wq/wk/wv, gate/up) in one launch; out[m][t*ldo[m] + i]. */
TB200_API int tb200_mul_mat_multi(int n, const tb200_weights *const *w, const float *x, long ldx, long tokens, float *const *out,
                                  const long *ldo);

/* The same in two steps, for callers that feed one activation tensor to several weight matrices (wq/wk/wv,
 * gate/up): quantize once into a resident activation object, then run any number of matmuls against it. */
typedef struct tb200_acts tb200_acts;
TB200_API tb200_acts *tb200_acts_create(int weight_type, long k, long max_tokens);
TB200_API int tb200_acts_quantize(tb200_acts *a, const float *x, long ldx, long tokens);
TB200_API int tb200_mul_mat_acts(const tb200_weights *w, const tb200_acts *a, float *out, long ldo);
TB200_API int tb200_mul_mat_acts_multi(int n, const tb200_weights *const *w, const tb200_acts *a, float *const *out, const long *ldo);
TB200_API void tb200_acts_free(tb200_acts *a);

/* Decode plan: a whole sequence of one-token matmul groups (a decode step: wq/wk/wv, wo, gate/up, down per layer, then the
 * vocabulary matrix) as ONE persistent kernel, one CTA per SM, the weight stream kept in flight across group boundaries by
 * per-warp TMA rings in shared memory (tb200_plan.cu). Group g starts when group g-1 is complete on every SM (stream-order
 * semantics: its input may be any function of earlier outputs). Each group first turns its fp32 input into the Q8_0
 * activations the reference's from_float would produce (ggml.c:12345-12373), optionally through one of the matmul's
 * neighbours in the reference graph (llama.cpp/src/llama.cpp:8099-8225 llm_build_norm / llm_build_ffn):
 *   TB200_PLAN_QUANT    x                                         -> quantize_row_q8_0 (bit-exact)
 *   TB200_PLAN_RMSNORM  x * 1/sqrt(mean(x^2) + eps) * norm_weight -> quantize   (ggml_rms_norm + ggml_mul)
 *   TB200_PLAN_SWIGLU   silu(x) * x2                              -> quantize   (ggml_silu + ggml_mul; x = gate, x2 = up)
 * and every matrix may add a residual on the way out (out = W.x + residual, ggml_add). All matrices of a plan Q4_0 (Q8_0
 * activations) or all Q4_K (the prologue then ends in quantize_row_q8_K, bit-exact likewise), K % 256 == 0, one token. Buffers must stay valid while the plan exists; weights, inputs and outputs may be shared between groups. */
enum { TB200_PLAN_QUANT = 0, TB200_PLAN_RMSNORM = 1, TB200_PLAN_SWIGLU = 2,
       /* OR-ed into `prologue`: the caller declares that this group reads nothing an earlier group of the same launch writes and
        * overwrites nothing one of them reads (independent matmuls: experts, branches, a micro-benchmark of one shape): it starts
        * without the grid barrier, so the weight stream of consecutive independent groups never stops */
       TB200_PLAN_INDEPENDENT = 0x100 };
typedef struct {
    int n;                                             /* matrices that share the input, <= TB200_GEMV_MAX_MATS */
    int prologue;                                      /* TB200_PLAN_* */
    const tb200_weights *w[TB200_GEMV_MAX_MATS];
    float *out[TB200_GEMV_MAX_MATS];                   /* fp32 [rows] */
    const float *residual[TB200_GEMV_MAX_MATS];        /* optional fp32 [rows] */
    const float *x;                                    /* fp32 [k], 16-byte aligned */
    const float *x2;                                   /* SWIGLU: fp32 [k] */
    const float *norm_weight;                          /* RMSNORM: fp32 [k] */
    float eps;                                         /* RMSNORM */
} tb200_plan_group;
typedef struct tb200_plan tb200_plan;
TB200_API tb200_plan *tb200_plan_create(int n_groups, const tb200_plan_group *groups);
TB200_API int tb200_plan_run(tb200_plan *p, int first, int count);   /* enqueue groups [first, first+count) as one launch */
TB200_API int tb200_plan_status(tb200_plan *p);   /* 0, or non-zero when a launch gave up waiting: 1 = grid barrier, 2 = a ring stage never landed, 3 = a peer's
                                                   * sequence number (multi-GPU). Sticky: from then on launches stop waiting (they drain at once, results undefined)
                                                   * until tb200_plan_reset */
TB200_API int tb200_plan_reset(tb200_plan *p);
/* debug: first call switches per-group phase stamps on (returns 0); later calls copy [n_groups][SMs][8] SM-clock stamps out */
TB200_API long tb200_plan_trace(tb200_plan *p, long long *out, long capacity);
TB200_API void tb200_plan_free(tb200_plan *p);
/* Multi-GPU decode plan, one process per GPU (declared after tb200_peers below): tb200_plan_create_peers. */

/* The matmul's neighbours for any token count (the decode plan fuses the same operations into its kernel):
 * fp32 device tensors, row-major with leading dimensions in floats. Same IEEE operations in the same order as ggml's CPU ops:
 *   tb200_rms_norm_mul_dev   y = x * 1/sqrtf(mean(x^2) + eps) * norm_weight     ggml_rms_norm (ggml.c:11850-11896) + ggml_mul
 *   tb200_silu_mul_dev       y = gate/(1 + expf(-gate)) * up                    ggml_silu (ggml.c:2393) + ggml_mul
 *   tb200_moe_combine_dev    out[t] = sum_j expert_out[t][j] * weights[t][j]    ggml_mul + ggml_add in expert order (llama.cpp:8303-8318)
 * and the first two straight into resident activations (transform, then the bit-exact from_float of the vec_dot type). */
TB200_API int tb200_rms_norm_mul_dev(const float *x, long ldx, const float *norm_weight, float eps, long tokens, long k, float *y, long ldy);
TB200_API int tb200_silu_mul_dev(const float *gate, const float *up, long ld, long tokens, long k, float *y, long ldy);
TB200_API int tb200_moe_combine_dev(const float *expert_out, const float *weights, long tokens, int used, long rows, float *out, long ldo);
/* the same graph ops one by one (what the ggml backend runs when a neighbour cannot be fused into a matmul: more than one token, or
 * an intermediate that something else reads): ggml_rms_norm, ggml_mul / ggml_add (ldb = 0: b is one row broadcast over the tokens,
 * ggml's broadcasting of a norm weight), ggml_silu. fp32, row-major, leading dimensions in floats; bit-for-bit ggml's CPU arithmetic
 * except expf in silu (see tb200_silu_mul_dev). */
TB200_API int tb200_rms_norm_dev(const float *x, long ldx, float eps, long tokens, long k, float *y, long ldy);
TB200_API int tb200_mul_dev(const float *a, long lda, const float *b, long ldb, long tokens, long k, float *y, long ldy);
TB200_API int tb200_add_dev(const float *a, long lda, const float *b, long ldb, long tokens, long k, float *y, long ldy);
TB200_API int tb200_silu_dev(const float *x, long ldx, long tokens, long k, float *y, long ldy);
TB200_API int tb200_acts_quantize_rmsnorm(tb200_acts *a, const float *x, long ldx, long tokens, const float *norm_weight, float eps, long k);
TB200_API int tb200_acts_quantize_swiglu(tb200_acts *a, const float *gate, const float *up, long ld, long tokens, long k);

/* Multi-GPU, one process per GPU: weights row-sharded (each rank uploads its row slice), the all-gather of the output
 * slices fused into the kernels as peer stores over NVLink (decode: the GEMV epilogue, one 64-byte segment per CTA and
 * peer; more than TB200_GEMV_MAX_TOKENS tokens: the tcgen05 GEMM epilogue, 128-byte segments, all four weight types).
 * The host shares one symmetric output buffer and one flag buffer ([TB200_MAX_PEERS] uint32, zeroed) per rank through CUDA
 * IPC handles (both from tb200_malloc). Matrix m's rows are this rank's slice starting at row row_lo[m] of a full output
 * with leading dimension ldo[m] stored at byte offset out_offset[m] of the symmetric buffer; every rank holds every
 * slice once tb200_peers_barrier() has passed. The reference moves such slices with cudaMemcpyPeerAsync + events per
 * device (llama.cpp/ggml/src/ggml-cuda.cu:1600-1676).
 * SLOT REUSE RULE: the barrier orders "all ranks have WRITTEN" before "any rank reads", not "all ranks have READ" before the
 * next write. A region of the symmetric buffer may therefore be written again only after a LATER barrier that every rank
 * enqueued after its last read of the region (e.g. give every matrix of a step its own offset, as bench.py does: a slot is
 * rewritten one step, i.e. >= 1 barrier, after its consumers ran; or alternate two offsets). A barrier that gives up waiting
 * (~2 s, a peer died) sets a sticky flag, tb200_peers_timed_out(): treat it as fatal, the epochs of the ranks no longer agree. */
typedef struct tb200_peers tb200_peers;
TB200_API int tb200_ipc_get_handle(const void *dev_ptr, void *handle64);
TB200_API void *tb200_ipc_open_handle(const void *handle64);
TB200_API int tb200_ipc_close_handle(void *p);
TB200_API tb200_peers *tb200_peers_create(int world, int rank, void *const *out_bases, void *const *flag_bases);
TB200_API void tb200_peers_free(tb200_peers *g);
TB200_API int tb200_mul_mat_acts_multi_gather(tb200_peers *g, int n, const tb200_weights *const *w, const tb200_acts *a,
                                              const size_t *out_offset, const long *ldo, const long *row_lo);
TB200_API int tb200_peers_barrier(tb200_peers *g);   /* enqueue: all ranks' slices of everything launched so far are visible after it */
TB200_API int tb200_peers_timed_out(tb200_peers *g);
/* The decode plan on row-sharded weights (north_star "sharded by output rows ... all-gather"): every rank creates
 * the SAME sequence of groups over its own row slices and runs it the same number of times. A matrix whose `out` pointer lies
 * inside this rank's symmetric buffer [out_bases[rank], + symmetric_bytes) is a row slice: the kernel stores every result at the
 * same offset of every rank's buffer (NVLink peer stores from the epilogue) and the barrier after such a group is grid-wide AND
 * cross-GPU - a sequence number published in the peers' flag buffers, words 8.. - so the next group, on every rank, reads the
 * full vector from its local buffer; the launch itself ends only after the last group's slices of all ranks have arrived. Groups
 * whose outputs are private (replicated weights) synchronise locally only. No collective call, no extra launch: the all-gather of
 * ggml-cuda.cu:1600-1676 is stores + two words. The SLOT REUSE RULE above applies to the offsets. */
TB200_API tb200_plan *tb200_plan_create_peers(tb200_peers *g, size_t symmetric_bytes, int n_groups, const tb200_plan_group *groups);
TB200_API int tb200_memset(void *dst, int value, size_t bytes);

/* GGUF tensor loader: tensors of a GGUF file (v2/v3) straight from an mmap to resident weights.
 * Mirrors the reader in llama.cpp/ggml/src/ggml.c:20896-21160 (gguf_init_from_file) for the parts this path needs:
 * header, metadata skipped except general.alignment, tensor infos, aligned data section. `data` pointers returned by
 * tb200_gguf_find() point into the mapping and stay valid until tb200_gguf_close(). tb200_gguf_upload() accepts tensors of every weight type above
 * (tb200_weights_upload's list) of rank 2 (a weight matrix; rows [row_lo, row_hi) when row_hi > row_lo = a rank's shard) or
 * rank 3 (ne[2] stacked expert matrices, the form tb200_mixmul_dev takes). */
typedef struct tb200_gguf tb200_gguf;
TB200_API tb200_gguf *tb200_gguf_open(const char *path);
TB200_API void tb200_gguf_close(tb200_gguf *g);
TB200_API long tb200_gguf_tensor_count(const tb200_gguf *g);
TB200_API const char *tb200_gguf_tensor_name(const tb200_gguf *g, long i);
TB200_API int tb200_gguf_find(const tb200_gguf *g, const char *name, int *type, int64_t ne[4], const void **data, size_t *bytes);
TB200_API tb200_weights *tb200_gguf_upload(const tb200_gguf *g, const char *name, long row_lo, long row_hi);

/* llamafile_sgemm on device pointers with pre-quantized activations B (GGUF blocks of Btype, ldb in
 * blocks). Same service matrix as llamafile_sgemm. */
TB200_API int tb200_sgemm_dev(const tb200_weights *w, long n, const void *B, long ldb, int Btype, float *C, long ldc);

/* Integer block dot products of one weight row set against one activation row, for parity checks:
 * out[i*nb + b] (Q4_0/Q8_0: int32 per 32-block; Q4_K: two int32 per super-block, as the oracle). */
TB200_API int tb200_blockdot_dev(const tb200_weights *w, const void *B, int Btype, int32_t *out);

/* llamafile_mixmul on device pointers. weights: `experts` consecutive resident matrices (one upload of
 * experts*rows rows). thought f32 [tokens][tasks][k] contiguous, plan i32 [tokens][thinkers],
 * result f32 [tokens][thinkers][rows]. Rows whose plan entry is outside [0, experts) are left untouched. */
TB200_API int tb200_mixmul_dev(const tb200_weights *w, int experts, const float *thought, int tasks,
                               const int32_t *plan, int thinkers, long tokens, float *result);

/* raw device memory helpers for hosts without a CUDA runtime of their own */
TB200_API void *tb200_malloc(size_t bytes);
TB200_API void tb200_free(void *p);
TB200_API int tb200_memcpy_h2d(void *dst, const void *src, size_t bytes);
TB200_API int tb200_memcpy_d2h(void *dst, const void *src, size_t bytes);
TB200_API void *tb200_host_alloc(size_t bytes);   /* pinned */
TB200_API void tb200_host_free(void *p);

/* counters: kernels launched by this module since load (bench.py's gpu_launches) */
TB200_API uint64_t tb200_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* TINYBLAS_B200_H */
