#!/bin/bash
# Builds ab/lib_<name>.so from the CURRENT sources with extra nvcc flags for battlespace.cu (A/B runs on the GPU box:
# BATTLESPACE_LIB=ab/lib_<name>.so).  The actor objects are reused from the last `make`.
# usage: profiles/build_variant.sh name [-DFLAG ...]
set -e
NAME=$1; shift
cd "$(dirname "$0")/../deep-rl-battlespace_b200/csrc"
mkdir -p ../../ab
nvcc -O3 -std=c++17 -lineinfo --fmad=false -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -Xcompiler -ffp-contract=off \
     "$@" -Xptxas -v -c -o /tmp/bs_$NAME.o battlespace.cu 2> /tmp/ptxas_$NAME.log
nvcc -gencode arch=compute_100a,code=sm_100a -cudart static -shared -o ../../ab/lib_$NAME.so /tmp/bs_$NAME.o actor_[1-8].o actor_dispatch.o
grep -A2 "rollout_kernelILi3ELb1" /tmp/ptxas_$NAME.log | tail -2
