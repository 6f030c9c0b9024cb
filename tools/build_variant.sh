#!/bin/bash
# tools/build_variant.sh NAME "-DFLAG=..."  ->  path-tracer_b200/ab/libpt_NAME.so (A/B builds of the kernels only;
# selected at run time with PT_B200_LIB=...; never loaded by default)
set -e
cd "$(dirname "$0")/../path-tracer_b200/csrc"
name=$1; shift
B=/tmp/ptb_variant_$name
mkdir -p $B ../ab
nvcc "$@" -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 --fmad=false -std=c++17 -Xcompiler -fPIC -Xptxas -v -c kernels.cu -o $B/kernels.o 2> $B/ptxas.log
grep -A3 "k_shadowI\|k_extendI\|k_shadeE" $B/ptxas.log | grep -E "registers|spill" | tr '\n' ' '; echo
objs=""
for o in lbvh render scene_host bvh_build isf_loader png jpeg profile_yaml synth json gltf_convert isf_binary; do objs="$objs _build/$o.o"; done
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../ab/libpt_$name.so $B/kernels.o $objs -lz -Xcompiler -pthread
echo "built ab/libpt_$name.so"
