#!/bin/bash
# Build liborbx variants that differ in ONE compile-time macro of ONE kernel file (measurements only).
# usage: tools/ab_variants.sh name:file:MACRO=value ...   -> orb_slam3_ros2_b200/liborbx_<name>.so
set -e
cd "$(dirname "$0")/../orb_slam3_ros2_b200/csrc"
make >/dev/null
ALL="orbx_api k_pyramid k_fast k_quadtree k_blur k_describe k_match k_guided k_stage k_synth k_bow"
for spec in "$@"; do
  IFS=: read name file macro <<< "$spec"
  nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -D$macro -c $file.cu -o build/${file}_$name.o
  objs=""
  for o in $ALL; do if [ $o = $file ]; then objs="$objs build/${file}_$name.o"; else objs="$objs build/$o.o"; fi; done
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../liborbx_$name.so $objs -lcudart -ldl
  echo built liborbx_$name.so
done
