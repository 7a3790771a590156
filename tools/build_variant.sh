#!/bin/bash
# A variant of the library with extra -D flags on kernels.cu: tools/build_variant.sh NAME -DPML_LB_WARPS=16 ...
# -> build/ab/NAME/lib.so (run with PML_LIB=build/ab/NAME/lib.so); the other objects come from photon_ml_b200/lib.
set -e
name=$1; shift
mkdir -p build/ab/$name
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -O2 --expt-relaxed-constexpr \
  "$@" -c photon_ml_b200/csrc/kernels.cu -o build/ab/$name/kernels.o
nvcc -shared -o build/ab/$name/lib.so build/ab/$name/kernels.o photon_ml_b200/lib/api.o photon_ml_b200/lib/pack.o \
  photon_ml_b200/lib/peer.o photon_ml_b200/lib/noalloc.o -gencode arch=compute_100a,code=sm_100a
echo built build/ab/$name/lib.so
