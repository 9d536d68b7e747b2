#!/bin/bash
# quick rebuild of graph.o only (+ link); usage: tools/rebuild_graph.sh [extra nvcc flags]
set -e
cd "$(dirname "$0")/../expann_b200/csrc"
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-Wall,-Wno-unused-function,-pthread -Xptxas -v --expt-relaxed-constexpr "$@" -c graph.cu -o build/graph.o 2> build/graph.ptxas.log || (cat build/graph.ptxas.log; false)
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libexpann_gpu.so build/api.o build/scan_f32.o build/merge_topk.o build/graph.o build/quant.o build/graph_build.o build/tc.o build/multi.o build/tc_i8.o build/tc_tf32.o -cudart static -Xcompiler -pthread
grep -A2 "graph_search_inline" build/graph.ptxas.log | grep "Used" | head -4
