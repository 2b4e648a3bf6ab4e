#!/bin/bash
# tools/build_variant.sh <tag> <nvcc -D flags...>: a tuned copy of the library under lib/variants/ (parameter sweeps)
set -e
tag=$1; shift
cd "$(dirname "$0")/../chaos-ray-tracing-course-2025_b200"
mkdir -p build/variants lib/variants
/usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -prec-div=true -prec-sqrt=true -ftz=false \
  -Xcompiler -fPIC,-O2,-ffp-contract=off -Xptxas -v "$@" -c csrc/crt_cuda.cu -o build/variants/crt_cuda_$tag.o 2> build/variants/ptxas_$tag.log
grep -A2 "k_wf_traceILi1ELb0" build/variants/ptxas_$tag.log | grep -E "registers" | sed "s/^/$tag: /"
objs=$(ls build/*.o | grep -v crt_cuda.o)
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o lib/variants/libcrt_b200_$tag.so build/variants/crt_cuda_$tag.o $objs -cudart static -lpthread
