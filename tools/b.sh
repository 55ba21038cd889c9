#!/bin/bash
# build libsrt_b200.so in-tree; fails loudly
set -e
cd "$(dirname "$0")/.."
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -prec-div=true -prec-sqrt=true -ftz=false \
  -Xcompiler -fPIC -shared -Iinclude "$@" sexy-raytracer_b200/csrc/srt_b200.cu -o sexy-raytracer_b200/libsrt_b200.so
echo "built $(date +%T) $(md5sum sexy-raytracer_b200/libsrt_b200.so | cut -c1-8)"
