#!/bin/bash
# profiles/build_variant.sh <name> <-D flags...>: a variant of libloner_b200.so with lb_mlp_umma.cu recompiled under extra
# -D switches (tuning A/B runs), written to profiles/variants/lib_<name>.so (git-ignored; swapped in on the GPU box only).
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p profiles/variants
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --expt-relaxed-constexpr -Xcompiler -fPIC \
     -Xcompiler -fvisibility=hidden -I include "$@" -c loner_b200/csrc/lb_mlp_umma.cu -o profiles/variants/umma_$name.o
objs=$(ls loner_b200/lib/obj/*.o | grep -v lb_mlp_umma.o)
nvcc -shared -gencode arch=compute_100a,code=sm_100a $objs profiles/variants/umma_$name.o -o profiles/variants/lib_$name.so -lcudart
echo profiles/variants/lib_$name.so
