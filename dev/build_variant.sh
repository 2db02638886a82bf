#!/bin/bash
# Experiment builds of the layer kernels only: dev/build_variant.sh NAME [-DFLAG ...] -> dev/variants/NAME.so
# (dev/kbench.py binds whatever symbols a build exports).  Not part of the product build.
set -e
cd "$(dirname "$0")/.."
name=$1; shift
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -shared --cudart static \
  -Xptxas -v "$@" -o dev/variants/$name.so curl_b200/csrc/curl_kernels.cu 2>&1 | grep -A2 "bwd_tma_kernelINS_7OpLayerELi1ELb1" | grep -E "registers|spill" | head -3
