#!/bin/bash
# A/B of the transform's resident blocks per SM: tools/_build/minb9 (-DGJD_TRANSFORM_MINB_ALT=9: 72 registers, quantisation
# words from global memory so that nine blocks fit the shared memory) as GJD_TRANSFORM_VARIANT=1 against the default (0).
out=gpurun_out; mkdir -p $out
V=tools/_build/minb9/libgjd.so
for i in 1 2; do GJD_LIB=$V timeout 200 python tools/stage_bench.py 3000 10 0,1 2>&1 | tail -2 >> $out/r01p_ab_minb9.txt; done
cat $out/r01p_ab_minb9.txt
GJD_LIB=$V GJD_TRANSFORM_VARIANT=1 timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -2 | tee -a $out/r01p_ab_minb9.txt
timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -2 | tee $out/r01p_pytest_default_after_quant.txt
