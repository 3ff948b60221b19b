#!/bin/bash
# Multi-GPU visit: bench.py --config c5 (24000 mixed images sharded by gjd_partition) and the e2e probe at N ranks.
# usage: tools/gpu_scale.sh <tag> <N> [probe] [c5] [c3]
tag=$1; N=$2; shift 2
out=gpurun_out; mkdir -p $out
nvidia-smi --query-gpu=name --format=csv,noheader | sort | uniq -c > $out/${tag}_n${N}_gpu.txt; nproc >> $out/${tag}_n${N}_gpu.txt
run() { if [ "$N" = 1 ]; then python "$@"; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 "$@"; fi; }
for what in "$@"; do
case $what in
probe) run tools/e2e_probe_multi.py 3000 > $out/${tag}_e2e_probe_n${N}.txt 2>$out/${tag}_e2e_probe_n${N}.err; cat $out/${tag}_e2e_probe_n${N}.txt ;;
c5) run bench.py --config c5 --gpus $N --steps 5 --warmup 3 > $out/${tag}_bench_c5_n${N}.json 2> $out/${tag}_bench_c5_n${N}.err; echo "c5 rc=$?"; cat $out/${tag}_bench_c5_n${N}.json; tail -3 $out/${tag}_bench_c5_n${N}.err ;;
c5free) GJD_BENCH_LOCKSTEP=0 GJD_BENCH_TRACE=1 run bench.py --config c5 --gpus $N --steps 5 --warmup 3 > $out/${tag}_bench_c5free_n${N}.json 2> $out/${tag}_bench_c5free_n${N}.err; echo "c5free rc=$?"; grep "^trace" $out/${tag}_bench_c5free_n${N}.err > $out/${tag}_c5free_trace_n${N}.txt; cat $out/${tag}_bench_c5free_n${N}.json | head -c 400 ;;
c5trace) GJD_BENCH_TRACE=1 run bench.py --config c5 --gpus $N --steps 5 --warmup 3 > $out/${tag}_bench_c5_n${N}.json 2> $out/${tag}_bench_c5_n${N}.err; echo "c5 rc=$?"; grep "^trace" $out/${tag}_bench_c5_n${N}.err > $out/${tag}_c5_trace_n${N}.txt; cat $out/${tag}_bench_c5_n${N}.json | head -c 400 ;;
c3) run bench.py --gpus $N --steps 10 --warmup 3 > $out/${tag}_bench_n${N}.json 2> $out/${tag}_bench_n${N}.err; echo "c3 rc=$?"; cat $out/${tag}_bench_n${N}.json; tail -3 $out/${tag}_bench_n${N}.err ;;
esac
done
