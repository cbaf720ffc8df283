#!/bin/bash
# profiles/build_variant.sh <name> <source.cu> <-D flags...>: a variant of libloner_b200.so with ONE source of loner_b200/csrc
# recompiled under extra -D switches (tuning A/B runs), written to profiles/variants/lib_<name>.so (git-ignored; swapped in
# on the GPU box only by profiles/run_variants.sh).  Needs an up-to-date `python -m loner_b200.build` for the other objects.
set -e
cd "$(dirname "$0")/.."
name=$1; src=$2; shift 2
mkdir -p profiles/variants
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --expt-relaxed-constexpr -Xcompiler -fPIC \
     -Xcompiler -fvisibility=hidden -I include "$@" -c loner_b200/csrc/$src -o profiles/variants/obj_$name.o
objs=$(ls loner_b200/lib/obj/*.o | grep -v "/${src%.cu}.o")
nvcc -shared -gencode arch=compute_100a,code=sm_100a $objs profiles/variants/obj_$name.o -o profiles/variants/lib_$name.so -lcudart
rm -f profiles/variants/obj_$name.o
echo profiles/variants/lib_$name.so
