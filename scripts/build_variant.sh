#!/bin/bash
# build_variant.sh NAME [-DFLAG=VALUE ...] -> ab/librad_b200_NAME.so : another build of the library for A/B timing
# (scripts/ab_build.py); ab/ is git-ignored and travels to the GPU box.
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p ab rad_b200/csrc/build
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-fvisibility=hidden "$@" \
     -c rad_b200/csrc/api.cu -o ab/api_$name.o
[ -f rad_b200/csrc/build/host_index.o ] || make -s -C rad_b200/csrc build/host_index.o
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ab/librad_b200_$name.so ab/api_$name.o rad_b200/csrc/build/host_index.o -cudart static -lpthread
rm -f ab/api_$name.o
echo built ab/librad_b200_$name.so
