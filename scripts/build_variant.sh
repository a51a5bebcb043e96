#!/bin/bash
# A/B build: scripts/build_variant.sh NAME [extra nvcc flags] -> instancesegmentation_b200/libisg_NAME.so
# (use with ISG_LIB=$PWD/instancesegmentation_b200/libisg_NAME.so)
set -e
cd "$(dirname "$0")/../instancesegmentation_b200"
name=$1; shift
mkdir -p build/$name
for s in isg_api engine_build engine_fk decode; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden \
    --expt-relaxed-constexpr "$@" -c csrc/$s.cu -o build/$name/$s.o &
done
wait
nvcc -shared -o libisg_$name.so build/$name/*.o
ls -la libisg_$name.so
