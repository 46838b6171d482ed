#!/bin/bash
# builds an experiment variant of the library: tools/build_variant.sh NAME -DHB_ZP=4 -DHB_XCH=2 ...  -> variants/libhb_NAME.so
set -e
cd "$(dirname "$0")/.."
NAME=$1; shift
mkdir -p variants
S=hyteg_b200/csrc
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared "$@" \
  $S/kernels.cu $S/stream_kernel.cu $S/shuffle_kernel.cu $S/runtime.cu $S/topology.cpp $S/stencil_assembly.cpp $S/solvers_capi.cpp \
  -o variants/libhb_$NAME.so
echo built $NAME
