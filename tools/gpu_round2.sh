#!/bin/bash
# Round-2 GPU-box visit.  usage: [NCU_SET=c3|420|422|dri] [NCU_KERNELS="..."] [AB_LIBS="default name..."] [AB_SET=..] tools/gpu_round2.sh <tag> [zz] [tests] [ablibs] [bench] [refarm] [c5] [configs] [launches] [ncu] [san] [ab] [micro]
tag=${1:-r02x}; shift
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,power.limit --format=csv > $out/${tag}_gpu.txt 2>&1
nproc >> $out/${tag}_gpu.txt; free -g | head -2 >> $out/${tag}_gpu.txt
for what in "$@"; do
case $what in
tests)
  timeout 1200 python -m pytest tests -m gpu -q > $out/${tag}_pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> $out/${tag}_pytest_gpu.txt
  tail -5 $out/${tag}_pytest_gpu.txt ;;
zz)
  # the allocation-free entry point alone, in a fresh process (it must not depend on what the other tests left behind)
  timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $out/${tag}_pytest_zz_alone.txt 2>&1; echo "zz alone rc=$?" >> $out/${tag}_pytest_zz_alone.txt
  tail -3 $out/${tag}_pytest_zz_alone.txt ;;
ablibs)
  bash tools/ab_libs.sh $tag ${AB_LIBS:-default} ;;
bench)
  timeout 600 python bench.py --steps 10 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"
  cat $out/${tag}_bench.json; tail -3 $out/${tag}_bench.err ;;
refarm)
  timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_reference_arm.json 2>> $out/${tag}_bench.err
  cat $out/${tag}_bench_reference_arm.json ;;
c5)
  timeout 900 python bench.py --config c5 --steps 5 --warmup 3 > $out/${tag}_bench_c5.json 2> $out/${tag}_bench_c5.err; echo "bench c5 rc=$?"
  cat $out/${tag}_bench_c5.json; tail -3 $out/${tag}_bench_c5.err ;;
configs)
  timeout 600 python tools/config_bench.py 0 > $out/${tag}_configs.txt 2>&1; cat $out/${tag}_configs.txt ;;
launches)
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv \
      python bench.py --steps 2 --warmup 1 --e2e-steps 1 --no-cpu-baseline > $out/${tag}_ncu_launch.log 2>&1; echo "launch list rc=$?" ;;
ncu)
  for k in ${NCU_KERNELS:-k_transform k_huff_write k_huff_sync}; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -f -o $out/${tag}_$k \
      env GJD_SUB_LOG2=12 python tools/profile_run.py 1000 2 ${NCU_SET:-c3} > $out/${tag}_ncu_$k.log 2>&1; echo "ncu $k rc=$?"
  done ;;
san)
  # compute-sanitizer over the small-image parity tests and the damaged-file fuzzer
  timeout 1500 compute-sanitizer --tool memcheck --print-limit 20 python -m pytest tests/test_gpu_parity.py -m gpu -x -q \
      -k "coefficients_bit_exact or pixels_bit_exact_batch or restart_intervals_bit_exact or grayscale or error_statuses or int16" \
      > $out/${tag}_sanitizer_memcheck_tests.txt 2>&1; echo "memcheck tests rc=$?" | tee -a $out/${tag}_sanitizer_memcheck_tests.txt
  timeout 900 compute-sanitizer --tool memcheck --print-limit 20 python tools/fuzz_gpu.py 200 7 > $out/${tag}_sanitizer_memcheck_fuzz.txt 2>&1
  echo "memcheck fuzz rc=$?" | tee -a $out/${tag}_sanitizer_memcheck_fuzz.txt
  timeout 1500 compute-sanitizer --tool racecheck --print-limit 20 python -m pytest tests/test_gpu_parity.py -m gpu -x -q \
      -k "pixels_bit_exact_batch or restart_intervals_bit_exact or grayscale_bit_exact" \
      > $out/${tag}_sanitizer_racecheck_tests.txt 2>&1; echo "racecheck rc=$?" | tee -a $out/${tag}_sanitizer_racecheck_tests.txt
  tail -4 $out/${tag}_sanitizer_*.txt ;;
ab)
  # A/B of the two transform kernels on the resident c3 batch (0 = k_transform, 2 = the cost-sorted k_transform2)
  timeout 600 python tools/stage_bench.py 3000 10 ${AB_VARIANTS:-0,2,0,2} > $out/${tag}_transform_ab.txt 2>&1; cat $out/${tag}_transform_ab.txt ;;
micro)
  ./tools/microbench/pipes > $out/${tag}_pipes.txt 2>&1; cat $out/${tag}_pipes.txt ;;
*) echo "unknown step $what" ;;
esac
done
ls -la $out | tail -30
