#!/bin/bash
# A/B builds of libgpe_b200.so with extra -D flags: build_variant.sh <tag> <flags...>  ->  ../gpe_b200/variants/libgpe_b200_<tag>.so
set -e
tag=$1; shift
NVF="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr"
mkdir -p /tmp/gpe_var_$tag ../gpe_b200/variants
nvcc $NVF "$@" -c jet_launch_tc.cu -o /tmp/gpe_var_$tag/jl.o &
nvcc $NVF "$@" -c capi.cu -o /tmp/gpe_var_$tag/capi.o &
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../gpe_b200/variants/libgpe_b200_$tag.so /tmp/gpe_var_$tag/capi.o build/point_ops.o \
    build/jet_launch_wp64.o build/jet_launch_wp128.o build/jet_launch_wp256.o /tmp/gpe_var_$tag/jl.o -cudart static
echo built $tag
