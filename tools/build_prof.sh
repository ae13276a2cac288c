#!/bin/bash
# variants/prof.so: the library with the small-N kernel's in-kernel phase clock (KC_GPS_PROFILE) compiled in
set -e
cd /root/repo
mkdir -p variants
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -diag-suppress 550 -DKC_GPS_PROFILE \
     -c kima_b200/csrc/kc_gp_small.cu -o variants/kc_gp_small_prof.o
objs=$(ls kima_b200/build/*.o | grep -v kc_gp_small.o)
nvcc -shared -o variants/prof.so $objs variants/kc_gp_small_prof.o -ldl
ls -la variants/prof.so
