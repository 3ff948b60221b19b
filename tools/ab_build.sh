#!/bin/bash
# Build an A/B variant of libgjd.so here (cross-compiled, no GPU needed): tools/ab_build.sh <name> <extra nvcc flags...>
#   -> tools/_build/<name>/libgjd.so ; run it on the box with GJD_LIB=tools/_build/<name>/libgjd.so
set -e
cd "$(dirname "$0")/.."
name=$1; shift
V=tools/_build/$name; mkdir -p $V
F="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-Wall,-Wno-unused-function -diag-suppress 177 $*"
nvcc $F -c gpu-jpeg-decoder_b200/csrc/gjd_kernels.cu -o $V/gjd_kernels.o &
nvcc $F -c gpu-jpeg-decoder_b200/csrc/gjd_api.cu -o $V/gjd_api.o &
g++ -O2 -std=c++17 -fPIC -c gpu-jpeg-decoder_b200/csrc/gjd_parse.cpp -o $V/gjd_parse.o
g++ -O2 -std=c++17 -fPIC -c gpu-jpeg-decoder_b200/csrc/gjd_writers.cpp -o $V/gjd_writers.o
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $V/libgjd.so $V/*.o -Xlinker --no-undefined -lpthread
echo "$V/libgjd.so: $*"
