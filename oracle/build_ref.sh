#!/usr/bin/env bash
# TEST INFRASTRUCTURE — builds oracle/_ref/libtb200_ref.so from the reference's OWN sources.
#
# Compiles, unmodified, the reference translation units that serve the hot path on x86:
#   llamafile/tinyblas_cpu_{sgemm,mixmul}_amd_{zen4,avx512f,avx2}.cpp, tinyblas_cpu_mixmul_amd_avx.cpp
#   (llamafile_mixmul_needs), llamafile/iqk_mul_mat_amd_{zen4,avx2}.cpp
# plus pristine llama.cpp/ggml/src/{ggml,ggml-quants,ggml-aarch64,ggml-alloc,ggml-backend}.c
# and links them with oracle/ref_wrap.cpp (our thin ctypes-friendly wrapper).
#
# The reference's own build (cosmocc + make setup) is not run. What `make setup` would do to the
# headers (llama.cpp.patches/apply-patches.sh:21-87: copy llamafile-files, renames.sh, patch -p0)
# is replayed in a scratch directory OUTSIDE the repo; no reference source is copied into the repo
# and the only outputs are under oracle/_ref/.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${TB200_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
S="${TB200_REF_SCRATCH:-${TMPDIR:-/tmp}/tb200_ref_build}"
JOBS="${JOBS:-8}"

if [ ! -d "$REF/llamafile" ]; then
  echo "build_ref: $REF not present; keeping prebuilt $OUT (if any)" >&2
  exit 0
fi
mkdir -p "$OUT"
STAMP="$OUT/.stamp"
SIG="$(cat "$HERE/ref_wrap.cpp" "$HERE/ref_ops.c" "$HERE/build_ref.sh" "$HERE/../include/tb200_ggml_backend.h" | sha1sum | cut -d' ' -f1)"
if [ -f "$OUT/libtb200_ref.so" ] && [ -f "$STAMP" ] && [ "$(cat "$STAMP")" = "$SIG" ]; then
  exit 0
fi

rm -rf "$S"; mkdir -p "$S/shim" "$S/obj"
cp -r "$REF/llama.cpp" "$S/llama.cpp"
cp -r "$REF/llamafile" "$S/llamafile"
chmod -R u+w "$S"
cp -r "$REF/llama.cpp.patches/llamafile-files/." "$S/llama.cpp/"
( cd "$S/llama.cpp" && bash "$REF/llama.cpp.patches/renames.sh" >/dev/null 2>&1 || true )
for p in ggml.c ggml.h ggml-quants.inc ggml-common.h ggml-backend.c ggml-backend.h \
         ggml-backend-impl.h ggml-alloc.c ggml-cuda.h ggml-metal.h; do
  ( cd "$S" && patch -p0 -s -N < "$REF/llama.cpp.patches/patches/$p.patch" )
done

# <cosmo.h> stand-in: the only Cosmopolitan bits the hot-path TUs use.
cat > "$S/shim/cosmo.h" <<'EOF'
#pragma once
#include <cpuid.h>
#include <stdint.h>
#include <stdlib.h>
#ifndef MIN
#define MIN(a, b) ((a) < (b) ? (a) : (b))
#endif
#ifndef MAX
#define MAX(a, b) ((a) > (b) ? (a) : (b))
#endif
#define bsr(x) (63 - __builtin_clzl(x))
#define X86_CHECK(F) 1
EOF

CXXF="-std=gnu++17 -O2 -fPIC -fvisibility=hidden -I$S -I$S/shim"
AVX2="-mtune=skylake -mavx -mf16c -mfma -mavx2"                     # llamafile/BUILD.mk:113,120-121
AVX512F="$AVX2 -mavx512f"                                           # BUILD.mk:124-125
ZEN4="$AVX2 -mavx512f -mavx512vl -mavx512vnni -mavx512bf16 -mavx512bw -mavx512dq"  # BUILD.mk:114,126-127
AVX="-mtune=sandybridge -mavx -mf16c"                               # BUILD.mk:116-117

pids=()
cc() { # name flags
  ( cd "$S" && g++ $CXXF $2 -c "llamafile/$1.cpp" -o "obj/$1.o" ) & pids+=($!)
}
cc tinyblas_cpu_sgemm_amd_zen4 "$ZEN4"
cc tinyblas_cpu_mixmul_amd_zen4 "$ZEN4"
cc iqk_mul_mat_amd_zen4 "$ZEN4"
cc tinyblas_cpu_sgemm_amd_avx512f "$AVX512F"
cc tinyblas_cpu_mixmul_amd_avx512f "$AVX512F"
cc tinyblas_cpu_sgemm_amd_avx2 "$AVX2"
cc tinyblas_cpu_mixmul_amd_avx2 "$AVX2"
cc iqk_mul_mat_amd_avx2 "$AVX2"
cc tinyblas_cpu_mixmul_amd_avx "$AVX"

# pristine upstream ggml (haswell ISA so the AVX2 quantize_row_q8_0 body, ggml-quants.c:940-990, is
# the one compiled and the .so also runs on hosts without AVX-512).
GG="$REF/llama.cpp/ggml"
# llamafile's own (patched) AVX2 build of ggml-quants: runtime quantizers + vec_dots, block_q8_K = {d,bsums,qs}
( cd "$S" && gcc -O2 -fPIC -fvisibility=hidden -mavx -mavx2 -mfma -mf16c -ffp-contract=off -DNDEBUG -w \
      -I. -Ishim -c llama.cpp/ggml-quants-amd-avx2.c -o obj/lf_quants_avx2.o ) & pids+=($!)
for f in ggml ggml-quants ggml-aarch64 ggml-alloc ggml-backend; do
  # default visibility: tests drive ggml's own graph / backend machinery (ggml_init, ggml_mul_mat, ggml_backend_graph_compute ...)
  # through ctypes to run the SAME graph on ggml's CPU backend and on the TB200 backend (tests/test_gpu_backend.py)
  ( gcc -O2 -fPIC -fvisibility=default -mavx -mavx2 -mfma -mf16c -ffp-contract=off -D_GNU_SOURCE \
        -DNDEBUG -w -Dquantize_row_q8_0=upstream_quantize_row_q8_0 -I"$GG/include" -I"$GG/src" -c "$GG/src/$f.c" -o "$S/obj/up_$f.o" ) & pids+=($!)
done
( g++ $CXXF $AVX2 -fopenmp -c "$HERE/ref_wrap.cpp" -o "$S/obj/ref_wrap.o" ) & pids+=($!)
# the matmul's neighbours (rms_norm, silu, mul, add) through the pristine ggml's own graph compute
( gcc -O2 -fPIC -fvisibility=hidden -w -I"$GG/include" -I"$GG/src" -I"$HERE/../include" -c "$HERE/ref_ops.c" -o "$S/obj/ref_ops.o" ) & pids+=($!)
for p in "${pids[@]}"; do wait "$p"; done

g++ -shared -fopenmp -o "$OUT/libtb200_ref.so" "$S"/obj/*.o -lm -lpthread
echo "$SIG" > "$STAMP"
rm -rf "$S"
echo "build_ref: wrote $OUT/libtb200_ref.so"
