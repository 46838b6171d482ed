#!/bin/bash
# TEST-ONLY host emulation build of the kernel sources (see hyteg_b200/csrc/hb_compat.h).
# Never loaded by the hyteg_b200 package; used by the CPU-only ("not gpu") tests for host logic.
set -e
cd "$(dirname "$0")/../.."
SRC=hyteg_b200/csrc
OUT=tests/emul/libhyteg_b200_emul.so
g++ -std=c++17 -O2 -fPIC -shared -DHB_EMUL -Wall -Wno-unused-function -Wno-unknown-pragmas \
  -x c++ $SRC/kernels.cu -x c++ $SRC/shuffle_kernel.cu -x c++ $SRC/p2_kernels.cu -x c++ $SRC/runtime.cu \
  -x c++ $SRC/topology.cpp $SRC/stencil_assembly.cpp $SRC/solvers_capi.cpp -o $OUT
echo built $OUT
