#!/bin/bash
# Tuning builds of wf_trace.cu: tools/build_variants.sh name "-DFLAG=.." [name "-D.."]...  -> tuning_libs/<name>.so (select with B2PATH_LIB)
set -e
cd "$(dirname "$0")/../mitsuba3_experiments_b200/csrc"
mkdir -p ../../tuning_libs
while [ $# -gt 1 ]; do
  name=$1; flags=$2; shift 2
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC -Xptxas -v $flags -c wf_trace.cu -o /tmp/wf_trace_$name.o 2> ../../tuning_libs/$name.ptxas.log
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../tuning_libs/$name.so capi.o render.o wavefront.o /tmp/wf_trace_$name.o wf_noalloc.o bvh_build.o tables.o -lcudart
  echo "$name: $(grep -A2 'k_wf_trace_packetILb0' ../../tuning_libs/$name.ptxas.log | grep -o 'Used [0-9]* registers\|[0-9]* bytes spill stores' | tr '\n' ' ')"
done
