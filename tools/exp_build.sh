#!/bin/bash
# usage: tools/exp_build.sh <name> <nvcc -D flags...>: a copy of the library with ntx_mlp.cu compiled under experiment flags
# (the other objects are reused) -> nerf_tex_b200/exp_<name>.so; run with NTX_LIB=nerf_tex_b200/exp_<name>.so
set -e
cd "$(dirname "$0")/.."
name=$1; shift
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden \
  --expt-relaxed-constexpr --extended-lambda "$@" -c nerf_tex_b200/csrc/ntx_mlp.cu -o /tmp/ntx_mlp_$name.o
objs=$(ls nerf_tex_b200/csrc/build/*.o | grep -v ntx_mlp.o)
nvcc -shared -o nerf_tex_b200/exp_$name.so /tmp/ntx_mlp_$name.o $objs -gencode arch=compute_100a,code=sm_100a
echo nerf_tex_b200/exp_$name.so
