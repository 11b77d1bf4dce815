#!/bin/bash
# Tuning variant of libmimgpu.so: recompiles ONE translation unit with extra -D flags and links it with the
# other objects of the regular build:  TAG=c3_256 TU=tu_admm.cu bash scripts/build_variant.sh -DMIMGPU_ADMM_C3_MAXT=256
# -> mim_solvers_b200/csrc/libmimgpu_<TAG>.so, selected at run time by MIMGPU_LIB=<path>.
set -eu
cd "$(dirname "$0")/../mim_solvers_b200/csrc"
TU=${TU:-tu_admm.cu}
O=${TU%.cu}_${TAG}.o
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xptxas -v "$@" -c -o $O $TU > build_${TAG}.log 2>&1
objs=$(ls mimgpu.o tu_small.o tu_stream.o tu_cta.o tu_admm.o tu_solve.o peaks.o noalloc.o | grep -v "^${TU%.cu}.o$")
nvcc -shared -o libmimgpu_${TAG}.so $objs $O
echo "built libmimgpu_${TAG}.so"
