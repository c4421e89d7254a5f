#!/bin/bash
# Build a variant of libjaco_b200.so (development tool).  usage: scripts/build_variant.sh <tag> [nvcc flags / -DNAME=VALUE ...]
#   -> mujoco_jaco_b200/libjaco_b200_<tag>.so ; flags starting with "+" replace the default fast-math flags (e.g. +ieee)
set -e
cd "$(dirname "$0")/../mujoco_jaco_b200"
tag=$1; shift
fm="-prec-div=false -prec-sqrt=true -ftz=true"
if [ "$1" = "+ieee" ]; then fm=""; shift; fi
mkdir -p build_var_$tag
for u in jaco_kernel_coop jaco_kernels; do
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 $fm -Xcompiler -fPIC "$@" -c -o build_var_$tag/$u.o csrc/$u.cu &
done
wait
/usr/local/cuda/bin/nvcc -shared -o libjaco_b200_$tag.so build_var_$tag/jaco_kernels.o build_var_$tag/jaco_kernel_coop.o
echo built $tag
