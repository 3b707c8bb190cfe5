#!/bin/bash
# Builds libcsm_b200.so (CUDA kernels for sm_100a behind the C ABI) and libcsm_host.so (the C++
# host layer) in-tree; one nvcc / g++ process per source file, in parallel (see Makefile).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
make -s -C "$HERE" -j"$(nproc)" all "$@"
