#!/bin/bash
# dev helper: build/libopal_<tag>.so for a list of "tag:-Dflags" variants of the fast kernels (run after `make -C opal_b200/csrc`)
set -e
cd "$(dirname "$0")/.."
mkdir -p build
if [ "$1" = "--clean" ]; then shift; find build -maxdepth 1 -type f \( -name "kf_*.o" -o -name "libopal_*.so" \) -delete; fi
C=opal_b200/csrc
NVCC=/usr/local/cuda/bin/nvcc
ARCH="-gencode arch=compute_100a,code=sm_100a"
for v in "$@"; do
  tag=${v%%:*}; flags=${v#*:}
  ( $NVCC $ARCH -ccbin /usr/bin/g++ -std=c++17 -O3 -lineinfo -Xcompiler -fPIC -DOPAL_KNS=fast -DOPAL_FAST=1 --fmad=true $flags -c $C/opal_kernels.cu -o build/kf_$tag.o &&
    $NVCC $ARCH -ccbin /usr/bin/g++ -shared -o build/libopal_$tag.so build/kf_$tag.o $C/kernels_strict.o $C/hgram.o $C/absorb.o $C/host.o -cudart static -ldl && echo built $tag ) &
done
wait
