#!/bin/bash
# after the in-CTA pre-merge: full -m gpu suite, then the small-shard op timings and launch
# lists again (compare with profiles/r01_small_shard_*), then the default bench
mkdir -p gpurun_out; rm -f gpurun_out/small2_*
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider -x > gpurun_out/pytest_gpu.log 2>&1; echo pytest=$?
tail -12 gpurun_out/pytest_gpu.log
for spec in "cfg5 8" "cfg4 8" "cfg1 1" "cfg5 1" "cfg2 1" "cfg3 1" "cfg4 1"; do
  set -- $spec
  python tools/kbench.py --configs $1 --tp $2 --ops v2 --iters 20 >> gpurun_out/small2_kbench.jsonl 2>> gpurun_out/small2_kbench.err
done
cat gpurun_out/small2_kbench.jsonl
for spec in "cfg5 8" "cfg4 8" "cfg1 1"; do
  set -- $spec
  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'paged_attention' -c 24 --csv \
    --log-file gpurun_out/small2_launches_$1_tp$2.csv python tools/kbench.py --configs $1 --tp $2 --ops v2 --iters 5 \
    > gpurun_out/small2_ncu_$1_tp$2.log 2>&1
  tail -4 gpurun_out/small2_launches_$1_tp$2.csv | awk -F'","' '{print $5, $NF}'
done
python bench.py > gpurun_out/bench_default.log 2> gpurun_out/bench_default.err; echo bench=$?
cat gpurun_out/bench_default.log
