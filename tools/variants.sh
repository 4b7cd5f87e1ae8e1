#!/bin/bash
# Experiment builds of libbvx.so with different -D switches: tools/variants.sh name "-DX=1 ..." [name2 "..."]
# -> 3dobject_detection_temporal_b200/libbvx_<name>.so  (run with BVX_LIB=<path>)
cd "$(dirname "$0")/../3dobject_detection_temporal_b200/csrc" || exit 1
while [ $# -ge 2 ]; do
  /usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC \
     --fmad=false -I../../include -I. $2 -shared -o ../libbvx_$1.so bvx.cu -lcudart || exit 1
  echo "built libbvx_$1.so ($2)"
  shift 2
done
