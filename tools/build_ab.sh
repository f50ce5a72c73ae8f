#!/bin/bash
# A/B builds of libb2r.so with different graph-kernel compile-time knobs: tools/build_ab.sh <tag> <nvcc -D flags...>
# -> poseestimator_b200/csrc/_ab/libb2r_<tag>.so (select with B2R_LIB=...)
set -e
cd "$(dirname "$0")/../poseestimator_b200/csrc"
tag=$1; shift
mkdir -p _ab/_obj_$tag
ARCH="-gencode arch=compute_100a,code=sm_100a"
objs=""
for f in api clique tail rt radix scale presteps nn_gemm refine raster graph graph_tc; do
  case $f in
    graph|graph_tc|rt|clique|tail) nvcc $ARCH -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -cudart static "$@" -c $f.cu -o _ab/_obj_$tag/$f.o & objs="$objs _ab/_obj_$tag/$f.o";;
    *) objs="$objs _obj/$f.o";;
  esac
done
wait
nvcc $ARCH -shared -cudart static -Xcompiler -pthread -o _ab/libb2r_$tag.so $objs
echo built _ab/libb2r_$tag.so
