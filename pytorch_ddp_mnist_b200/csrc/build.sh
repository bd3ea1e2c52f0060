#!/bin/bash
# Builds libddpmnist_b200.so in-tree for sm_100a (cross-compiles without a GPU).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="${DDPM_OUT:-${HERE}/../libddpmnist_b200.so}"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
"${NVCC}" -shared -Xcompiler -fPIC -O3 -std=c++17 -lineinfo \
    -gencode arch=compute_100a,code=sm_100a \
    ${DDPM_PTXAS_V:+-Xptxas -v} ${DDPM_DEFS:-} \
    -o "${OUT}" "${HERE}/mnist_kernels.cu" "${HERE}/persist_kernel.cu" "${HERE}/fc1_tc.cu" "${HERE}/convnet_kernels.cu" "${HERE}/convnet_tc.cu" "${HERE}/cdf5_reader.cpp" ${DDPM_EXTRA_SRC:-} -lcudart
echo "built ${OUT}"
