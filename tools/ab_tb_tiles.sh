#!/bin/bash
# A/B of the transform block's tile count on the resident C3 batch (GJD_TB_TILES_MAX), two rounds, alternating.
out=gpurun_out; mkdir -p $out
for i in 1 2; do for t in 32 8 16 64 128; do
  echo -n "tb_tiles $t round $i: " >> $out/r01p_ab_tb_tiles.txt
  GJD_TB_TILES_MAX=$t timeout 200 python tools/stage_bench.py 3000 10 2>&1 | tail -1 >> $out/r01p_ab_tb_tiles.txt
done; done
cat $out/r01p_ab_tb_tiles.txt
GJD_TB_TILES_MAX=64 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
