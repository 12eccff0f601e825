#!/bin/bash
# build libt4c.so with extra -D flags into tesseract_b200/csrc/variants/libt4c_<name>.so (experiments: T4C_LIB=<path>)
# usage: tools/build_variant.sh <name> [-DFLAG ...]
set -e
name=$1; shift
cd "$(dirname "$0")/.."
out=tesseract_b200/csrc/variants; mkdir -p $out
F="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC"
nvcc $F "$@" -c tesseract_b200/csrc/t4c.cu -o $out/$name.t4c.o &
nvcc $F "$@" -fmad=false -c tesseract_b200/csrc/t4c_exact.cu -o $out/$name.exact.o &
wait
nvcc -shared -gencode arch=compute_100a,code=sm_100a $out/$name.t4c.o $out/$name.exact.o -o $out/libt4c_$name.so
rm -f $out/$name.t4c.o $out/$name.exact.o
ls -la $out/libt4c_$name.so
