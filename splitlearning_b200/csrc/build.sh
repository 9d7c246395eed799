#!/bin/sh
# Builds splitlearning_b200/libsplitnn.so for sm_100a, in-tree (the .so travels with gpurun).
set -e
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=${SN_OUT:-"$HERE/../libsplitnn.so"}
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
$NVCC -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
  -Xcompiler -fPIC -shared ${SN_NVCC_EXTRA} \
  -o "$OUT" "$HERE/splitnn.cu" "$HERE/lda_partition.cpp" -lcudart
echo "built $OUT"
