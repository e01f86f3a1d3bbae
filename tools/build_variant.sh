#!/bin/bash
# build an alternative librt2cuda into gpurun_variants/<name>.so with extra nvcc flags: ./tools/build_variant.sh name -DFOO ...
name=$1; shift
mkdir -p gpurun_variants
CSRC=raytracer2_b200/csrc
/usr/local/cuda/bin/nvcc "$@" -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -Xcompiler -fPIC,-ffp-contract=off -shared \
  -o gpurun_variants/$name.so $CSRC/rt2_api.cu $CSRC/rt2_scene.cu $CSRC/rt2_build.cu $CSRC/rt2_render.cu $CSRC/rt2_noalloc.cu $CSRC/rt2_group.cu $CSRC/rt2_passes*.cu 2>&1 | grep -E "error" -A3
ls -la gpurun_variants/$name.so
