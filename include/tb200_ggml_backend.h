/* tinyBLAS-B200 as a ggml backend: MUL_MAT / MUL_MAT_ID with device-resident weights and activations.
 *
 * llamafile's GPU module is a ggml backend living in a DSO that cannot link against the host binary: llamafile/cuda.c
 * dlopen()s it, imports a 12-symbol table (llamafile/cuda.c:104-120, 701-753) and hands it the host's ggml helpers as a
 * table of function pointers, `struct ggml_backend_api` (llama.cpp.patches/patches/ggml-backend-impl.h.patch:20-58). The
 * backend itself is an instance of `struct ggml_backend` whose interface table is `struct ggml_backend_i`
 * (llama.cpp/ggml/src/ggml-backend-impl.h:86-131; the CUDA instance: llama.cpp/ggml/src/ggml-cuda.cu:2976-2997).
 *
 * This header declares layout-compatible mirrors of exactly those interface structures (this repository never includes
 * the reference's headers; tests/test_abi.py compiles a probe against the reference's own headers and compares every
 * offset and enum value used here) and the module's exports, named after the symbols cuda.c imports:
 *
 *   ggml_cuda_link                           -> tb200_backend_link
 *   ggml_backend_cuda_init                   -> tb200_backend_init
 *   ggml_backend_cuda_buffer_type            -> tb200_backend_buffer_type
 *   ggml_backend_cuda_host_buffer_type       -> tb200_backend_host_buffer_type
 *   ggml_backend_cuda_split_buffer_type      -> tb200_backend_split_buffer_type   (NULL: row splits go through the peer API)
 *   ggml_backend_cuda_reg_devices            -> tb200_backend_reg_devices
 *   ggml_backend_cuda_get_device_count       -> tb200_backend_get_device_count
 *   ggml_backend_cuda_get_device_memory      -> tb200_backend_get_device_memory
 *   ggml_backend_cuda_get_device_description -> tb200_backend_get_device_description
 *   ggml_backend_cuda_get_device_properties  -> tb200_backend_get_device_properties
 *   ggml_backend_cuda_register_host_buffer   -> tb200_backend_register_host_buffer
 *   ggml_backend_cuda_unregister_host_buffer -> tb200_backend_unregister_host_buffer
 *
 * What the backend computes: GGML_OP_MUL_MAT and GGML_OP_MUL_MAT_ID for the weight types of include/tinyblas_b200.h, the view ops
 * (no-ops) and the matmul's neighbours in llm_build_norm / llm_build_ffn on fp32 matrices - RMS_NORM, MUL (same shape or a row
 * broadcast: the norm weight), ADD, UNARY(SILU) - so that a transformer block's activations stay on the device; supports_op()
 * answers false for everything else (rope, soft_max, the batched attention matmuls, ...), which ggml's scheduler keeps on another
 * backend. At one token those neighbours are folded into the decode plan when nothing else reads the intermediates: RMSNorm * weight
 * and silu(gate) * up become a matmul group's prologue, a residual add its epilogue. Weights are repacked once, on the device, the first time a graph uses them; a run of consecutive
 * one-token Q4_0 MUL_MAT nodes is executed as ONE persistent kernel (the decode plan, tb200_plan.cu), cached per graph the
 * way ggml-cuda caches its CUDA graph (ggml-cuda.cu:2467-2700). */
#ifndef TB200_GGML_BACKEND_H
#define TB200_GGML_BACKEND_H

#include "tinyblas_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- enum values used (llama.cpp/ggml/include/ggml.h:443-528 enum ggml_op, :331-336 enum ggml_status, ggml-backend.h:29-33) */
enum {
    TB200_GGML_OP_NONE = 0, TB200_GGML_OP_MUL_MAT = 23, TB200_GGML_OP_MUL_MAT_ID = 24,
    TB200_GGML_OP_RESHAPE = 30, TB200_GGML_OP_VIEW = 31, TB200_GGML_OP_PERMUTE = 32, TB200_GGML_OP_TRANSPOSE = 33,
    /* the matmul's neighbours in llm_build_norm / llm_build_ffn: taken so that activations stay on the device, and
     * fused into the decode plan where the graph allows */
    TB200_GGML_OP_ADD = 2, TB200_GGML_OP_MUL = 6, TB200_GGML_OP_RMS_NORM = 20, TB200_GGML_OP_UNARY = 63,
    TB200_GGML_UNARY_OP_SILU = 10,          /* enum ggml_unary_op, kept in op_params[0] of a GGML_OP_UNARY node */
    TB200_GGML_TENSOR_FLAG_OUTPUT = 2,      /* enum ggml_tensor_flag */
};
enum { TB200_GGML_STATUS_ALLOC_FAILED = -2, TB200_GGML_STATUS_FAILED = -1, TB200_GGML_STATUS_SUCCESS = 0, TB200_GGML_STATUS_ABORTED = 1 };
enum { TB200_GGML_BUFFER_USAGE_ANY = 0, TB200_GGML_BUFFER_USAGE_WEIGHTS = 1, TB200_GGML_BUFFER_USAGE_COMPUTE = 2 };

/* struct ggml_tensor up to `data` is tb200_ggml_tensor_view (tinyblas_b200.h); `name` and `extra` follow and are not read,
 * so the view holds for GGML_MAX_NAME = 64 (upstream) and 128 (llamafile, ggml.h.patch:28-29) alike. */

/* struct ggml_cgraph (ggml.h:655-667): only the node list is read */
typedef struct {
    int size, n_nodes, n_leafs;
    struct ggml_tensor **nodes;
    struct ggml_tensor **grads_;
    struct ggml_tensor **leafs_;
    /* struct ggml_hash_set visited_hash_set; enum ggml_cgraph_eval_order order;  (not read) */
} tb200_ggml_cgraph_view;

typedef struct tb200_ggml_backend_buffer_type *tb200_buft_t;
typedef struct tb200_ggml_backend_buffer *tb200_buffer_t;
typedef struct tb200_ggml_backend *tb200_backend_t;

/* ggml-backend-impl.h:18-37 */
struct tb200_ggml_backend_buffer_type_i {
    const char *(TB200_CALL *get_name)(tb200_buft_t buft);
    tb200_buffer_t (TB200_CALL *alloc_buffer)(tb200_buft_t buft, size_t size);
    size_t (TB200_CALL *get_alignment)(tb200_buft_t buft);
    size_t (TB200_CALL *get_max_size)(tb200_buft_t buft);
    size_t (TB200_CALL *get_alloc_size)(tb200_buft_t buft, const struct ggml_tensor *tensor);
    bool (TB200_CALL *is_host)(tb200_buft_t buft);
};
struct tb200_ggml_backend_buffer_type {
    struct tb200_ggml_backend_buffer_type_i iface;
    void *context;
};

/* ggml-backend-impl.h:42-60 */
struct tb200_ggml_backend_buffer_i {
    const char *(TB200_CALL *get_name)(tb200_buffer_t buffer);
    void (TB200_CALL *free_buffer)(tb200_buffer_t buffer);
    void *(TB200_CALL *get_base)(tb200_buffer_t buffer);
    void (TB200_CALL *init_tensor)(tb200_buffer_t buffer, struct ggml_tensor *tensor);
    void (TB200_CALL *set_tensor)(tb200_buffer_t buffer, struct ggml_tensor *tensor, const void *data, size_t offset, size_t size);
    void (TB200_CALL *get_tensor)(tb200_buffer_t buffer, const struct ggml_tensor *tensor, void *data, size_t offset, size_t size);
    bool (TB200_CALL *cpy_tensor)(tb200_buffer_t buffer, const struct ggml_tensor *src, struct ggml_tensor *dst);
    void (TB200_CALL *clear)(tb200_buffer_t buffer, uint8_t value);
    void (TB200_CALL *reset)(tb200_buffer_t buffer);
};
struct tb200_ggml_backend_buffer {
    struct tb200_ggml_backend_buffer_i iface;
    tb200_buft_t buft;
    void *context;
    size_t size;
    int usage;
};

/* ggml-backend-impl.h:86-131 */
struct tb200_ggml_backend_i {
    const char *(TB200_CALL *get_name)(tb200_backend_t backend);
    void (TB200_CALL *free)(tb200_backend_t backend);
    tb200_buft_t (TB200_CALL *get_default_buffer_type)(tb200_backend_t backend);
    void (TB200_CALL *set_tensor_async)(tb200_backend_t backend, struct ggml_tensor *tensor, const void *data, size_t offset, size_t size);
    void (TB200_CALL *get_tensor_async)(tb200_backend_t backend, const struct ggml_tensor *tensor, void *data, size_t offset, size_t size);
    bool (TB200_CALL *cpy_tensor_async)(tb200_backend_t backend_src, tb200_backend_t backend_dst, const struct ggml_tensor *src, struct ggml_tensor *dst);
    void (TB200_CALL *synchronize)(tb200_backend_t backend);
    void *(TB200_CALL *graph_plan_create)(tb200_backend_t backend, const struct ggml_cgraph *cgraph);
    void (TB200_CALL *graph_plan_free)(tb200_backend_t backend, void *plan);
    void (TB200_CALL *graph_plan_update)(tb200_backend_t backend, void *plan, const struct ggml_cgraph *cgraph);
    int (TB200_CALL *graph_plan_compute)(tb200_backend_t backend, void *plan);
    int (TB200_CALL *graph_compute)(tb200_backend_t backend, struct ggml_cgraph *cgraph);
    bool (TB200_CALL *supports_op)(tb200_backend_t backend, const struct ggml_tensor *op);
    bool (TB200_CALL *supports_buft)(tb200_backend_t backend, tb200_buft_t buft);
    bool (TB200_CALL *offload_op)(tb200_backend_t backend, const struct ggml_tensor *op);
    void *(TB200_CALL *event_new)(tb200_backend_t backend);
    void (TB200_CALL *event_free)(void *event);
    void (TB200_CALL *event_record)(void *event);
    void (TB200_CALL *event_wait)(tb200_backend_t backend, void *event);
    void (TB200_CALL *event_synchronize)(void *event);
};
struct tb200_ggml_backend {
    const uint8_t *guid;   /* ggml_guid_t = uint8_t (*)[16] */
    struct tb200_ggml_backend_i iface;
    void *context;
};

/* The host's services, member for member llamafile's `struct ggml_backend_api` (ggml-backend-impl.h.patch:20-58). The module
 * calls only the ones marked [used]; the rest keep the table layout. */
typedef tb200_backend_t (TB200_CALL *tb200_backend_init_fn)(const char *params, void *user_data);
struct tb200_ggml_backend_api {
    bool *FLAG_log_disable;
    void (TB200_CALL *exit)(int);
    void (TB200_CALL *free)(void *);                                                     /* [used] */
    void *(TB200_CALL *malloc)(size_t);                                                  /* [used] */
    char *(TB200_CALL *getenv)(const char *);                                            /* [used] */
    long (TB200_CALL *write)(int, const void *, long);                                   /* [used] log sink */
    void (TB200_CALL *ggml_backend_register)(const char *, tb200_backend_init_fn, tb200_buft_t, void *);   /* [used] */
    tb200_buffer_t (TB200_CALL *ggml_backend_buffer_init)(tb200_buft_t, struct tb200_ggml_backend_buffer_i, void *, size_t);   /* [used] */
    tb200_buffer_t (TB200_CALL *ggml_backend_cpu_buffer_from_ptr)(void *, size_t);       /* [used] pinned host buffers */
    tb200_buft_t (TB200_CALL *ggml_backend_cpu_buffer_type)(void);
    size_t (TB200_CALL *ggml_backend_buft_get_alloc_size)(tb200_buft_t, struct ggml_tensor *);
    tb200_buffer_t (TB200_CALL *ggml_backend_buft_alloc_buffer)(tb200_buft_t, size_t);
    bool (TB200_CALL *ggml_backend_is_cpu)(tb200_backend_t);
    void (TB200_CALL *ggml_backend_tensor_get)(const struct ggml_tensor *, void *, size_t, size_t);
    void (TB200_CALL *ggml_backend_tensor_set)(struct ggml_tensor *, const void *, size_t, size_t);
    bool (TB200_CALL *ggml_is_quantized)(int);
    size_t (TB200_CALL *ggml_type_size)(int);
    int64_t (TB200_CALL *ggml_blck_size)(int);
    bool (TB200_CALL *ggml_is_transposed)(const struct ggml_tensor *);
    size_t (TB200_CALL *ggml_nbytes)(const struct ggml_tensor *);                        /* [used] */
    int (TB200_CALL *ggml_get_unary_op)(const struct ggml_tensor *);
    int64_t (TB200_CALL *ggml_nelements)(const struct ggml_tensor *);
    int64_t (TB200_CALL *ggml_nrows)(const struct ggml_tensor *);
    bool (TB200_CALL *ggml_is_permuted)(const struct ggml_tensor *);
    bool (TB200_CALL *ggml_is_contiguous)(const struct ggml_tensor *);
    const char *(TB200_CALL *ggml_op_name)(int);
    const char *(TB200_CALL *ggml_type_name)(int);
    size_t (TB200_CALL *ggml_element_size)(const struct ggml_tensor *);
    size_t (TB200_CALL *ggml_row_size)(int, int64_t);
    void (TB200_CALL *ggml_rope_yarn_corr_dims)(int, int, float, float, float, float[2]);
    const char *(TB200_CALL *ggml_op_desc)(const struct ggml_tensor *);
    bool (TB200_CALL *ggml_backend_buffer_is_host)(tb200_buffer_t);
    bool (TB200_CALL *ggml_guid_matches)(const uint8_t *, const uint8_t *);
    bool (TB200_CALL *ggml_is_empty)(const struct ggml_tensor *);
    int (TB200_CALL *ggml_backend_buffer_get_usage)(tb200_buffer_t);
    bool (TB200_CALL *ggml_are_same_shape)(const struct ggml_tensor *, const struct ggml_tensor *);
    bool (TB200_CALL *ggml_is_contiguous_1)(const struct ggml_tensor *);
    bool (TB200_CALL *ggml_is_contiguous_2)(const struct ggml_tensor *);
};

/* llamafile's `struct ggml_cuda_device_properties` (ggml-cuda.h.patch:18-24) */
struct tb200_device_properties {
    char name[256];
    size_t totalGlobalMem;
    int multiProcessorCount;
    int major, minor;
    char compute[8];
};

/* ---- exports (see the table at the top) */
TB200_API bool tb200_backend_link(const struct tb200_ggml_backend_api *api);   /* true iff an sm_100 device is usable */
TB200_API tb200_backend_t tb200_backend_init(int device);
TB200_API bool tb200_backend_is_tb200(tb200_backend_t backend);
TB200_API tb200_buft_t tb200_backend_buffer_type(int device);
TB200_API tb200_buft_t tb200_backend_host_buffer_type(void);                   /* pinned host memory */
TB200_API tb200_buft_t tb200_backend_split_buffer_type(const float *tensor_split);
TB200_API int tb200_backend_reg_devices(void);
TB200_API int tb200_backend_get_device_count(void);
TB200_API void tb200_backend_get_device_memory(int device, size_t *free_bytes, size_t *total_bytes);
TB200_API void tb200_backend_get_device_description(int device, char *description, size_t description_size);
TB200_API void tb200_backend_get_device_properties(int device, struct tb200_device_properties *properties);
TB200_API bool tb200_backend_register_host_buffer(void *buffer, size_t size);
TB200_API void tb200_backend_unregister_host_buffer(void *buffer);
/* statistics of the last graph_compute of a backend: kernels launched, and nodes that went through the decode plan */
TB200_API void tb200_backend_last_graph_stats(tb200_backend_t backend, int *launches, int *plan_nodes);

#ifdef __cplusplus
}
#endif
#endif /* TB200_GGML_BACKEND_H */
