#!/bin/bash
# Development build of the N = 0 executor only into build_ab/lib_<tag>.so (linked against the in-tree objects of the other translation
# units).  Usage: tools/octbuild.sh <tag> [extra nvcc flags, e.g. -DVIEKF_OCTET_MINB=4]
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
TAG=$1; shift
mkdir -p $ROOT/build_ab
NVCC=/usr/local/cuda/bin/nvcc
ARCH="-gencode arch=compute_100a,code=sm_100a"
C=$ROOT/vi-ekf_b200/csrc
$NVCC -O3 -std=c++17 -lineinfo $ARCH -Xcompiler -fPIC -Xptxas -v "$@" -c $C/viekf_frag.cu -o $ROOT/build_ab/frag_$TAG.o 2> $ROOT/build_ab/ptxas_frag_$TAG.log || { grep -i error $ROOT/build_ab/ptxas_frag_$TAG.log | head -20; exit 1; }
$NVCC $ARCH -shared -o $ROOT/build_ab/lib_$TAG.so $ROOT/build_ab/frag_$TAG.o $C/viekf_kernel.o $C/viekf_tile.o $C/viekf_octet.o $C/viekf_capi.o -lcudart
grep -A2 "frag_kernel" $ROOT/build_ab/ptxas_frag_$TAG.log | grep -v Compiling | tr '\n' ' '; echo
