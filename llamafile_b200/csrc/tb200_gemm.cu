// Prefill path placeholder — replaced by the tcgen05 GEMM.
#include "tb200_common.cuh"
bool tb200_gemm_available() { return false; }
int tb200_launch_gemm(tb200_ctx *, const tb200_weights *, long, long, const int8_t *, const int2 *, const void *, long, long, float *, long) {
    return TB200_ERR_UNSUPPORTED;
}
void tb200_gemm_release(tb200_weights *) {}
