/* Plain-C loader for libtinyblas_b200.so, written the way llamafile/cuda.c:694-753 loads its GPU DSO:
 * dlopen(RTLD_LAZY) -> dlsym every entry point into a table -> link() handshake that reports whether
 * a usable device exists. A llamafile maintainer would compile this file into the host binary (with
 * cosmo_dlopen/cosmo_dlsym in place of dlopen/dlsym) and route llamafile_sgemm()/llamafile_mixmul()
 * through the table when tb200_loader_ready() is true; see INTEGRATION.md. */
#include <dlfcn.h>
#include <stdio.h>
#include <stdlib.h>
#include <unistd.h>

#include "tinyblas_b200.h"

struct tb200_table {
    bool (*link)(const struct tb200_host_api *);
    bool (*sgemm)(long, long, long, const void *, long, const void *, long, void *, long, int, int, int, int, int);
    bool (*mixmul)(const struct ggml_compute_params *, const struct ggml_tensor *, const struct ggml_tensor *,
                   const struct ggml_tensor *, struct ggml_tensor *);
    size_t (*mixmul_needs)(const struct ggml_tensor *, const struct ggml_tensor *, const struct ggml_tensor *);
    bool (*mixmul_iqk)(long, long, long, int, int, const void *, const void *, float *, long, long, const void *, int, int);
    bool (*quantize_row_q8_0)(const float *, void *, int64_t);
    bool (*quantize_row_q8_K)(const float *, void *, int64_t);
    const char *(*last_error)(void);
};

static struct tb200_table g_tb;
static void *g_lib;
static int g_ready;

static long host_write(int fd, const void *buf, size_t n) { return (long)write(fd, buf, n); }
static const char *host_getenv(const char *k) { return getenv(k); }

#define IMPORT(field, name)                                   \
    do {                                                      \
        *(void **)(&g_tb.field) = dlsym(g_lib, name);         \
        if (!g_tb.field) {                                    \
            fprintf(stderr, "tb200_loader: missing %s\n", name); \
            ok = 0;                                           \
        }                                                     \
    } while (0)

__attribute__((visibility("default"))) int tb200_loader_init(const char *dso_path) {
    if (g_ready) return 1;
    g_lib = dlopen(dso_path, RTLD_LAZY);
    if (!g_lib) {
        fprintf(stderr, "tb200_loader: %s\n", dlerror());
        return 0;
    }
    int ok = 1;
    IMPORT(link, "tb200_link");
    IMPORT(sgemm, "llamafile_sgemm");
    IMPORT(mixmul, "llamafile_mixmul");
    IMPORT(mixmul_needs, "llamafile_mixmul_needs");
    IMPORT(mixmul_iqk, "llamafile_mixmul_iqk");
    IMPORT(quantize_row_q8_0, "tb200_quantize_row_q8_0");
    IMPORT(quantize_row_q8_K, "tb200_quantize_row_q8_K");
    IMPORT(last_error, "tb200_last_error");
    if (!ok) {
        dlclose(g_lib);
        g_lib = NULL;
        return 0;
    }
    static const struct tb200_host_api api = {malloc, free, host_write, host_getenv};
    g_ready = g_tb.link(&api) ? 1 : 0;   /* false: no sm_100 device -> host keeps its CPU path */
    return g_ready;
}

__attribute__((visibility("default"))) int tb200_loader_ready(void) { return g_ready; }

/* trampolines with the reference signatures: what llamafile/sgemm.cpp:128-145 would call */
__attribute__((visibility("default"))) bool tb200_loader_sgemm(long m, long n, long k, const void *A, long lda, const void *B,
                                                              long ldb, void *C, long ldc, int ith, int nth, int Atype,
                                                              int Btype, int Ctype) {
    return g_ready && g_tb.sgemm(m, n, k, A, lda, B, ldb, C, ldc, ith, nth, Atype, Btype, Ctype);
}
__attribute__((visibility("default"))) bool tb200_loader_mixmul(const struct ggml_compute_params *p, const struct ggml_tensor *w,
                                                               const struct ggml_tensor *x, const struct ggml_tensor *plan,
                                                               struct ggml_tensor *out) {
    return g_ready && g_tb.mixmul(p, w, x, plan, out);
}
