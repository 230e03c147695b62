#!/bin/sh
# Instrumented (-DFB200_TRACE) build of the library for tools/attn_trace.py / gemm_trace.py; never used by the product path.
set -e
cd "$(dirname "$0")/../funasr_b200"
mkdir -p build/trace
for f in api frontend rowwise gemm_sm100 attention_sm100 cif_decode conformer relpos_attention_sm100 bicif runtime stack_sm100; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden \
    -DFB200_TRACE -c csrc/$f.cu -o build/trace/$f.o &
done
wait
nvcc -shared -o libfunasr_b200_trace.so build/trace/*.o -gencode arch=compute_100a,code=sm_100a
