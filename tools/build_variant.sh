#!/bin/bash
# usage: tools/build_variant.sh NAME [-DFOO=1 ...]  — builds build/variants/libb200pt_NAME.so with extra nvcc flags (kernel tuning only;
# select it at run time with B200PT_LIB=build/variants/libb200pt_NAME.so).  build/ is git-ignored but ships to the GPU box.
set -e
name=$1; shift
root=$(cd "$(dirname "$0")/.." && pwd)
src=$root/optixpathtracer_b200/csrc
out=$root/build/variants/$name
mkdir -p "$out"
flags="-gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=default $*"
( cd "$src"
  nvcc $flags -c bvh_build.cu -o "$out/bvh_build.o" &
  nvcc $flags -c trace.cu -o "$out/trace.o" &
  nvcc $flags -fmad=${SHADE_FMAD:-false} -c shade.cu -o "$out/shade.o" &
  nvcc $flags -c api.cu -o "$out/api.o" &
  wait )
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o "$root/build/variants/libb200pt_$name.so" "$out"/*.o -lcudart -ldl
rm -rf "$out"
echo "built build/variants/libb200pt_$name.so"
