#!/bin/bash
# tools/build_variant.sh <tag> <nvcc -D flags...>: a tuned copy of the library under lib/variants/ (parameter sweeps;
# bench / tools pick it up through $CRT_B200_LIB).  Both device translation units are rebuilt with the flags.
set -e
tag=$1; shift
cd "$(dirname "$0")/../chaos-ray-tracing-course-2025_b200"
mkdir -p build/variants lib/variants
FLAGS="-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -prec-div=true -prec-sqrt=true -ftz=false -Xcompiler -fPIC,-O2,-ffp-contract=off -Xptxas -v"
/usr/local/cuda/bin/nvcc $FLAGS "$@" -c csrc/crt_cuda.cu -o build/variants/crt_cuda_$tag.o 2> build/variants/ptxas_$tag.log &
/usr/local/cuda/bin/nvcc $FLAGS "$@" -c csrc/crt_gi.cu -o build/variants/crt_gi_$tag.o 2> build/variants/ptxas_gi_$tag.log &
wait
for k in "k_wf_trace_leanILb0" "k_render_persistentILb0ELb0ELb1ELb0"; do
  grep -A2 "Function properties for _ZN6crtdev[0-9]*$k" build/variants/ptxas_$tag.log | grep -E "spill|registers" | sed "s/^/$tag $k: /"
done
objs=$(ls build/*.o | grep -v "crt_cuda.o\|crt_gi.o")
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o lib/variants/libcrt_b200_$tag.so build/variants/crt_cuda_$tag.o build/variants/crt_gi_$tag.o $objs -cudart static -lpthread
