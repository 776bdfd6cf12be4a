#!/bin/bash
# Build a variant of the library that differs only in the trace kernels' translation unit (lj_wf_trace.cu), e.g.
#   tools/build_trace_variant.sh noprmt -DLJ_SOME_KNOB=1      -> lajolla_public_b200/variants/lib_noprmt.so
# and run it with LJ_LIB=lajolla_public_b200/variants/lib_noprmt.so (tools/profile_stages.py, tests).
set -e
name=$1; shift
cd "$(dirname "$0")/../lajolla_public_b200"
mkdir -p variants build/var
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -Xcompiler -fPIC -Xptxas -v -rdc=false "$@" \
    -c csrc/lj_wf_trace.cu -o build/var/lj_wf_trace_$name.o 2> build/var/lj_wf_trace_$name.ptxas.log
objs=$(ls build/obj/*.o | grep -v lj_wf_trace.o)
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o variants/lib_$name.so $objs build/var/lj_wf_trace_$name.o -cudart static -lz
grep -A2 "k_wf_extendILb0\|k_wf_shadowILb0" build/var/lj_wf_trace_$name.ptxas.log | grep "Used\|spill"
