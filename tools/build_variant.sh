#!/bin/bash
# build a variant of libmpcb.so with extra nvcc defines: tools/build_variant.sh <name> <flags...>  ->  mpc_motion_planner_b200/libmpcb_<name>.so
name=$1; shift
cd "$(dirname "$0")/.."
mkdir -p build
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Iinclude -Impc_motion_planner_b200/csrc -Xcompiler -fPIC "$@" -c mpc_motion_planner_b200/csrc/mpcb_kernels.cu -o build/var_${name}.o 2>&1 | grep -i "error"
nvcc -shared -o mpc_motion_planner_b200/libmpcb_${name}.so build/var_${name}.o build/mpcb_capi.o build/mpcb_planner_host.o -cudart shared -Xlinker -rpath=/usr/local/cuda/lib64
