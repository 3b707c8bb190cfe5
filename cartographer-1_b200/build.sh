#!/bin/bash
# Builds, in-tree and for sm_100a only:
#   libcsm_b200.so  the C ABI (include/csm_b200.h): CUDA kernels + thin host glue
#   libcsm_host.so  the C++ host layer with the reference's class interfaces
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
$NVCC -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
  -Xcompiler -fPIC -shared -I"$HERE/../include" -I"$HERE/csrc" \
  -o "$HERE/libcsm_b200.so" "$HERE"/csrc/*.cu "$@"
g++ -O3 -DNDEBUG -std=c++17 -fPIC -ffp-contract=off -Wall -pthread -shared \
  -I"$HERE/../include" -I"$HERE/host" \
  -o "$HERE/libcsm_host.so" \
  $(find "$HERE/host" -name '*.cc') \
  -L"$HERE" -lcsm_b200 -Wl,-rpath,'$ORIGIN'
