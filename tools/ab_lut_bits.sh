#!/bin/bash
out=gpurun_out; mkdir -p $out
V=tools/_build/lut9/libgjd.so
for i in 1 2; do
  timeout 200 python tools/stage_bench.py 3000 10 > $out/r01p_ab_lut10_$i.txt 2>&1; tail -1 $out/r01p_ab_lut10_$i.txt
  GJD_LIB=$V timeout 200 python tools/stage_bench.py 3000 10 > $out/r01p_ab_lut9_$i.txt 2>&1; tail -1 $out/r01p_ab_lut9_$i.txt
done
GJD_LIB=$V timeout 500 python -m pytest tests -m gpu -x -q > $out/r01p_pytest_lut9.log 2>&1; echo "rc=$?" >> $out/r01p_pytest_lut9.log; tail -3 $out/r01p_pytest_lut9.log
GJD_LIB=$V timeout 200 python tools/config_bench.py > $out/r01p_configs_lut9.txt 2>&1; tail -5 $out/r01p_configs_lut9.txt
timeout 200 python tools/config_bench.py > $out/r01p_configs_lut10.txt 2>&1; tail -5 $out/r01p_configs_lut10.txt
