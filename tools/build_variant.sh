#!/bin/bash
# Build the whole library with extra -D flags into build_variants/lib_<name>.so (A/B runs of launch shapes and code
# options on the GPU box; tools/run_variants.sh copies them over the library in turn).
# Usage: tools/build_variant.sh <name> [-DFLAG=VALUE ...]
name=$1; shift
mkdir -p build_variants
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared "$@" \
  -o build_variants/lib_$name.so minicfd_b200/csrc/libmcfd.cu
