#!/bin/bash
# dev aid: scripts/build_variant.sh <name> <source.cu> "<extra nvcc -D flags>" -> transformerengine_b200/_build/variants/libte_<name>.so
# (the named source recompiled with the flags, every other object taken from the regular build)
set -e
name=$1; src=$2; flags=$3
B=transformerengine_b200/_build; V=$B/variants; mkdir -p $V
/usr/local/cuda/bin/nvcc -std=c++17 -O3 -lineinfo --expt-relaxed-constexpr -gencode arch=compute_100a,code=sm_100a \
  -Xcompiler -fPIC $flags -x cu -c transformerengine_b200/csrc/$src -o $V/$name.o
objs=$(ls $B/*.o | grep -v "/$src.o")
/usr/local/cuda/bin/nvcc -shared -o $V/libte_$name.so $objs $V/$name.o -gencode arch=compute_100a,code=sm_100a \
  -Xlinker --version-script=transformerengine_b200/csrc/exports.map -ldl
echo $V/libte_$name.so
