#!/bin/bash
# A/B builds of libb2r.so with different graph-kernel compile-time knobs: tools/build_ab.sh <tag> <nvcc -D flags...>
# -> poseestimator_b200/csrc/_ab/libb2r_<tag>.so (select with B2R_LIB=...)
set -e
cd "$(dirname "$0")/../poseestimator_b200/csrc"
tag=$1; shift
mkdir -p _ab/_obj_$tag
ARCH="-gencode arch=compute_100a,code=sm_100a"
nvcc $ARCH -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -cudart static "$@" -c graph.cu -o _ab/_obj_$tag/graph.o
nvcc $ARCH -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -cudart static "$@" -c graph_tc.cu -o _ab/_obj_$tag/graph_tc.o
objs=""
for f in api clique rt radix scale presteps nn_gemm refine raster; do objs="$objs _obj/$f.o"; done
nvcc $ARCH -shared -cudart static -Xcompiler -pthread -o _ab/libb2r_$tag.so _ab/_obj_$tag/graph.o _ab/_obj_$tag/graph_tc.o $objs
echo built _ab/libb2r_$tag.so
