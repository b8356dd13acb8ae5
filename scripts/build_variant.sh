#!/bin/bash
# usage: build_variant.sh name [extra nvcc flags]  -> gym-ctr-reach_b200/variants/lib_<name>.so
set -e
cd "$(dirname "$0")/../gym-ctr-reach_b200"; mkdir -p variants
name=$1; shift
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC -shared -cudart static "$@" -o variants/lib_$name.so csrc/ctr_kernels.cu
echo built variants/lib_$name.so
