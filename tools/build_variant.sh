#!/bin/bash
# Builds an experimental variant of libtinyblas_b200.so that differs in ONE translation unit's -D flags, for A/B runs on the
# GPU box (TB200_LIB=tools/_variants/<name>.so python tools/plan_probe.py ...). The regular build must exist (objects reused).
#   tools/build_variant.sh <name> <file.cu> [-DX=Y ...]
set -e
cd "$(dirname "$0")/.."
name=$1; src=$2; shift 2
mkdir -p tools/_variants
obj=tools/_variants/${name}.o
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden --expt-relaxed-constexpr "$@" -c llamafile_b200/csrc/$src -o $obj
others=$(ls llamafile_b200/_obj/*.o | grep -v "/${src%.cu}.o")
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o tools/_variants/${name}.so $others $obj -Xcompiler -fPIC -lpthread -ldl
rm -f $obj
echo tools/_variants/${name}.so
