#!/bin/bash
# usage: build_variant.sh name [-Dflags...]   -> variants/name.so (kernel tuning / instrumented builds)
name=$1; shift
mkdir -p variants
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC \
  --expt-relaxed-constexpr -shared "$@" -o variants/$name.so rotors_mpc_controller_b200/csrc/nmpc_abi.cu -lcudart
