#!/bin/bash
# usage: scripts/build_variant.sh NAME [-DFLAG ...]  -> build_variants/libavtubes_NAME.so (dev experiments)
set -e
cd "$(dirname "$0")/.."
mkdir -p build_variants
name=$1; shift
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared "$@" \
  -o build_variants/libavtubes_$name.so audio_visual_tubes_b200/csrc/*.cu -lcudart -lcuda
