#!/bin/bash
# Development build of the tile executor only (configs[3] instantiations, -DVIEKF_DEV_FAST) into build_ab/lib_<tag>.so, linked
# against the in-tree objects of the other two translation units.  Usage: tools/devbuild.sh <tag> [extra nvcc flags]
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
TAG=$1; shift
mkdir -p $ROOT/build_ab
NVCC=/usr/local/cuda/bin/nvcc
ARCH="-gencode arch=compute_100a,code=sm_100a"
$NVCC -O3 -std=c++17 -lineinfo $ARCH -Xcompiler -fPIC -DVIEKF_DEV_FAST -Xptxas -v "$@" -c $ROOT/vi-ekf_b200/csrc/viekf_tile.cu -o $ROOT/build_ab/tile_$TAG.o 2> $ROOT/build_ab/ptxas_$TAG.log || { grep -i error $ROOT/build_ab/ptxas_$TAG.log | head -20; exit 1; }
$NVCC $ARCH -shared -o $ROOT/build_ab/lib_$TAG.so $ROOT/build_ab/tile_$TAG.o $ROOT/vi-ekf_b200/csrc/viekf_kernel.o $ROOT/vi-ekf_b200/csrc/viekf_octet.o $ROOT/vi-ekf_b200/csrc/viekf_frag.o $ROOT/vi-ekf_b200/csrc/viekf_capi.o -lcudart
grep -A2 "Function properties" $ROOT/build_ab/ptxas_$TAG.log | grep -v "^--" | paste - - - | sed 's/ptxas info    : Function properties for //' | c++filt | sed 's/_INTERNAL_[0-9a-f_]*viekf_tile_cu_[0-9a-f]*:://g' | sed 's/ptxas info    ://' | cut -c1-230 | head -${LINES_OUT:-10}
