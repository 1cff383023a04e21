"""tinyBLAS-B200: B200-native (sm_100a) quantized matmul behind llamafile's llamafile_sgemm /
llamafile_mixmul entry points. The product is libtinyblas_b200.so (C ABI, include/tinyblas_b200.h);
this package is the thin Python host layer used by tests and bench.py: it loads the library with
ctypes and mirrors the reference's interface. There is no CPU fallback: without the CUDA library or
without a B200 every compute call raises."""
from .lib import (acts_quantize_rmsnorm, acts_quantize_swiglu, moe_combine, rms_norm_mul, silu_mul, rms_norm, mul, add, silu, backend_init, backend_last_graph_stats, backend_link, PLAN_QUANT, PLAN_RMSNORM, PLAN_SWIGLU, PLAN_INDEPENDENT, Plan, Acts, GgufFile, mul_mat_acts, mul_mat_multi, mul_mat_acts_multi, F16, F32, I32, IQ4_XS, Q4_0, Q4_1, Q4_K, Q5_0, Q5_K, Q6_K, Q8_0, Q8_1, Q8_K, TB200Error, Weights, blockdot, device_count, launch_count, lib,
                  link, mixmul, mul_mat, mul_mat_noalloc, quantize_q8_0, quantize_q8_K, set_stream, sgemm_dev, sm_count, synchronize)
from .sgemm import (GgmlComputeParams, GgmlTensor, llamafile_mixmul, llamafile_mixmul_iqk, llamafile_mixmul_needs,
                    llamafile_sgemm, fp32_to_fp16_row, quantize_row_q8_0, quantize_row_q8_K, register_weights, unregister_weights, weights_cache_bytes, weights_cache_clear)

__all__ = [n for n in dir() if not n.startswith("_")]
