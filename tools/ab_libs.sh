#!/bin/bash
# A/B of prebuilt variant libraries (tools/ab_build.sh) on the resident c3 batch, two alternating rounds.
# usage: [AB_SET=c5 AB_IMAGES=n] tools/ab_libs.sh <tag> <name>...     ("default" = the shipped build)
tag=$1; shift
out=gpurun_out; mkdir -p $out
{
for i in 1 2; do
  for v in "$@"; do
    if [ "$v" = default ]; then lib=gpu-jpeg-decoder_b200/build/libgjd.so; else lib=tools/_build/$v/libgjd.so; fi
    echo -n "round $i  $v: "; GJD_LIB=$lib timeout 300 python tools/stage_bench.py ${AB_IMAGES:-3000} 10 0 ${AB_SET:-c3} 2>&1 | tail -1
  done
done
} > $out/${tag}_ab_libs.txt 2>&1
cat $out/${tag}_ab_libs.txt
