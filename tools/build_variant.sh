#!/bin/bash
# tools/build_variant.sh <name> <source.cu> [nvcc flags...]: variants/<name>.so = the library with kc_gp_small.o replaced
set -e
cd /root/repo
name=$1; src=$2; shift 2
mkdir -p variants
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -diag-suppress 550 -Ikima_b200/csrc "$@" \
     -c $src -o variants/$name.o
objs=$(ls kima_b200/build/*.o | grep -v kc_gp_small.o)
nvcc -shared -o variants/$name.so $objs variants/$name.o -ldl 2>/dev/null
rm variants/$name.o
ls variants/$name.so
