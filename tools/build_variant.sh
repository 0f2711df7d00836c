#!/bin/bash
# tools/build_variant.sh NAME [nvcc -D flags...]: a variant build of both libraries under build_variants/NAME (experiments only)
set -e
cd "$(dirname "$0")/../lapis-silo_b200"
name=$1; shift
out=../build_variants/$name
mkdir -p $out
nvcc "$@" -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared -o $out/libsilo_b200.so csrc/silo_b200.cu
g++ -std=c++17 -O2 -fPIC -Wno-free-nonheap-object -Wno-alloc-size-larger-than -shared -o $out/libsilo_host.so host/capi_db.cpp host/capi_engine.cpp host/capi_synth.cpp -L$out -lsilo_b200 -Wl,-rpath,'$ORIGIN' -ldl -lpthread
