#!/bin/bash
# scratch/build_variant.sh name [-DMACRO=v ...]  ->  scratch/libptb_<name>.so built from the working tree with extra defines
name=$1; shift
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -shared "$@" -o scratch/libptb_$name.so pathtracerpython_b200/csrc/ptb_api.cu
