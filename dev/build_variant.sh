#!/bin/bash
# Experiment builds of the layer kernels only: dev/build_variant.sh NAME [-DFLAG ...] -> dev/variants/NAME.so
# (dev/kbench.py binds whatever symbols a build exports).  Not part of the product build.
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p dev/variants
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -shared --cudart static \
  -Xptxas -v "$@" -o dev/variants/$name.so curl_b200/csrc/curl_kernels.cu 2>&1 | grep -A2 -E "bwd_tma_kernelINS_7OpLayerELi1ELb1|fwd_tma_kernelINS_7OpLayerELi1E|fwd_kernelINS_7OpLayerELi1ELi4" | grep -E "registers|spill" | tr '\n' ' '; echo
