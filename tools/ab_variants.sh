#!/bin/bash
# Build liborbx variants that differ in compile-time macros of ONE kernel file (measurements only).
# usage: tools/ab_variants.sh name:file:MACRO=value[,MACRO=value...] ...   -> orb_slam3_ros2_b200/liborbx_<name>.so
# run one with ORBX_LIB=orb_slam3_ros2_b200/liborbx_<name>.so python bench.py ...
set -e
cd "$(dirname "$0")/../orb_slam3_ros2_b200/csrc"
make >/dev/null
ALL=$(ls *.cu | sed 's/\.cu$//')
for spec in "$@"; do
  IFS=: read name file macros <<< "$spec"
  defs=""; for m in ${macros//,/ }; do defs="$defs -D$m"; done
  nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -Xptxas -v $defs -c $file.cu -o build/${file}_$name.o 2> build/${file}_$name.ptxas.log
  objs=""
  for o in $ALL; do if [ $o = $file ]; then objs="$objs build/${file}_$name.o"; else objs="$objs build/$o.o"; fi; done
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../liborbx_$name.so $objs -lcudart -ldl
  echo built liborbx_$name.so
done
