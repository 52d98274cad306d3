#!/bin/bash
# Builds librt_b200 variants with different -D experiment switches (kernels only differ): tools/build_variants.sh "tag:-DA=1 -DB=0" ...
# Select one at run time with RT_B200_LIB=<path>.  Experiment tooling, not part of the product build.
set -eu
PKG=triangles_sdf_cpu_raytracing_b200
OBJ=$PKG/csrc/build
OUT=$PKG/variants
mkdir -p $OUT
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC"
make -s $PKG/librt_b200.so
for spec in "$@"; do
  tag=${spec%%:*}; defs=${spec#*:}
  (
    $NV $defs -fmad=false -c $PKG/csrc/rt_kernels_exact.cu -o $OUT/exact_$tag.o
    $NV $defs -fmad=true -c $PKG/csrc/rt_kernels_fast.cu -o $OUT/fast_$tag.o
    $NV -shared -o $OUT/librt_b200_$tag.so $OUT/exact_$tag.o $OUT/fast_$tag.o $OBJ/rt_api.o $OBJ/bvh_builder.o $OBJ/bvh_gpu.o $OBJ/ingest.o $OBJ/rt_host.o -cudart shared
    rm -f $OUT/exact_$tag.o $OUT/fast_$tag.o
    echo "built $OUT/librt_b200_$tag.so ($defs)"
  ) &
done
wait
