#!/bin/bash
# Build a variant of libjaco_b200.so that differs only in the hybrid-kernel unit (jaco_kernel_coop.cu).
# usage: scripts/build_variant.sh <tag> [-DNAME=VALUE ...]   ->  mujoco_jaco_b200/libjaco_b200_<tag>.so
set -e
cd "$(dirname "$0")/../mujoco_jaco_b200"
tag=$1; shift
mkdir -p build_var_$tag
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -prec-div=false -prec-sqrt=false -ftz=true -Xcompiler -fPIC "$@" -c -o build_var_$tag/jaco_kernel_coop.o csrc/jaco_kernel_coop.cu
/usr/local/cuda/bin/nvcc -shared -o libjaco_b200_$tag.so build/jaco_kernels.o build_var_$tag/jaco_kernel_coop.o
echo built $tag
