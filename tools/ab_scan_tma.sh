#!/bin/bash
# A/B of the TMA-staged scan ring (BASELINE.json north_star: "compressed scans are staged into shared memory by TMA bulk
# copies") against the shipped register window that reads global memory one word ahead (GjdWin).
#   build (here, no GPU needed):  tools/ab_scan_tma.sh build     -> tools/_build/tma/libgjd.so  (-DGJD_SCAN_TMA=1)
#   run (GPU box):                tools/ab_scan_tma.sh run <tag> -> gpurun_out/<tag>_tma_scan_ab.txt
set -e
cd "$(dirname "$0")/.."
V=tools/_build/tma
if [ "$1" = build ]; then
  mkdir -p $V
  F="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-Wall,-Wno-unused-function -diag-suppress 177 -DGJD_SCAN_TMA=1 ${GJD_TMA_EXTRA}"
  nvcc $F -c gpu-jpeg-decoder_b200/csrc/gjd_kernels.cu -o $V/gjd_kernels.o
  nvcc $F -c gpu-jpeg-decoder_b200/csrc/gjd_api.cu -o $V/gjd_api.o
  g++ -O2 -std=c++17 -fPIC -c gpu-jpeg-decoder_b200/csrc/gjd_parse.cpp -o $V/gjd_parse.o
  g++ -O2 -std=c++17 -fPIC -c gpu-jpeg-decoder_b200/csrc/gjd_writers.cpp -o $V/gjd_writers.o
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $V/libgjd.so $V/*.o -Xlinker --no-undefined -lpthread
  echo "SASS of the variant:"; cuobjdump -sass $V/libgjd.so | grep -oE "UBLKCP[.A-Z0-9]*|SYNCS[.A-Z0-9]*" | sort | uniq -c
  exit 0
fi
tag=${2:-r02x}; out=gpurun_out; mkdir -p $out
{
echo "== SASS of the TMA build (cuobjdump -sass $V/libgjd.so | grep UBLKCP/SYNCS) =="
cuobjdump -sass $V/libgjd.so | grep -oE "UBLKCP[.A-Z0-9]*|SYNCS[.A-Z0-9]*" | sort | uniq -c
echo "== parity of the TMA build (pytest -m gpu, bit-exact tests) =="
GJD_LIB=$V/libgjd.so timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "bit_exact or synthetic_seeded or subsequence_lengths or huffman_stress or restart" 2>&1 | tail -3
for i in 1 2; do
  echo "== round $i: shipped build (register window over global memory) =="
  timeout 300 python tools/stage_bench.py 3000 10 0 2>&1 | tail -1
  echo "== round $i: TMA scan ring (cp.async.bulk + mbarrier per thread, ${GJD_RING_CHUNK:-32}-byte chunks, double-buffered) =="
  GJD_LIB=$V/libgjd.so timeout 300 python tools/stage_bench.py 3000 10 0 2>&1 | tail -1
done
} > $out/${tag}_tma_scan_ab.txt 2>&1
cat $out/${tag}_tma_scan_ab.txt
