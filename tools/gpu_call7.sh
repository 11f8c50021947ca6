#!/bin/bash
# Where does the time go on the small shards?  Event-timed op (no profiler) first, then the ncu
# launch list of the same command: per-kernel durations of the main and the merge kernel
# (serialised and cold under ncu, so read them as shares / orders of magnitude).
mkdir -p gpurun_out
for spec in "cfg5 8" "cfg4 8" "cfg1 1" "cfg5 1"; do
  set -- $spec
  python tools/kbench.py --configs $1 --tp $2 --ops v2 --iters 20 >> gpurun_out/small_kbench.jsonl 2>> gpurun_out/small_kbench.err; echo kbench_$1_tp$2=$?
done
cat gpurun_out/small_kbench.jsonl
for spec in "cfg5 8" "cfg4 8" "cfg1 1" "cfg5 1"; do
  set -- $spec
  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'paged_attention' -c 24 --csv \
    --log-file gpurun_out/small_launches_$1_tp$2.csv python tools/kbench.py --configs $1 --tp $2 --ops v2 --iters 5 \
    > gpurun_out/small_ncu_$1_tp$2.log 2>&1; echo ncu_$1_tp$2=$?
  tail -6 gpurun_out/small_launches_$1_tp$2.csv
done
