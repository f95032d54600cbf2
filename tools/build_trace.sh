#!/bin/bash
# Developer tool: build/libtrace.so = the library with the GEMV per-CTA timeline compiled in (-DB200_MMVQ_TRACE), for tools/trace_mmvq.py
set -e
cd "$(dirname "$0")/.."
python -m koboldcpp_b200.build
mkdir -p build/trace
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden -I include --expt-relaxed-constexpr \
  -DB200_MMVQ_TRACE -x cu -c koboldcpp_b200/csrc/mmvq.cu -o build/trace/mmvq.o
objs=$(ls build/obj/*.o | grep -v '/mmvq.o')
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build/libtrace.so $objs build/trace/mmvq.o -lcuda
echo built build/libtrace.so
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden -I include --expt-relaxed-constexpr \
  -DB200_FA_TRACE -x cu -c koboldcpp_b200/csrc/fattn.cu -o build/trace/fattn.o
objs=$(ls build/obj/*.o | grep -v '/mmvq.o' | grep -v '/fattn.o')
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build/libtrace.so $objs build/trace/mmvq.o build/trace/fattn.o -lcuda
echo built build/libtrace.so with the attention timeline
