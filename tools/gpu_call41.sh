#!/bin/bash
# cfg2 (SURVEY 8d's headline target shape): bench line, ncu launch list and ncu --set full of the stream kernel.
mkdir -p gpurun_out
python bench.py --config cfg2 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r01end_bench_cfg2_n1.json 2> gpurun_out/r01end_cfg2.err; echo bench=$?
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'paged_attention|reshape_and_cache' -c 40 --csv \
  --log-file gpurun_out/r01end_cfg2_launches.csv python bench.py --config cfg2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01end_cfg2_ncu_list.log 2>&1; echo list=$?
tail -4 gpurun_out/r01end_cfg2_launches.csv | awk -F'","' '{print $5, $NF}'
ncu --set full --clock-control none --import-source on -k regex:'paged_attention_stream_kernel' -s 4 -c 1 -f \
  -o gpurun_out/r01end_cfg2_stream_kernel python bench.py --config cfg2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01end_cfg2_ncu_full.log 2>&1; echo full=$?
python -c "
import json; r=json.load(open('gpurun_out/r01end_bench_cfg2_n1.json')); print(r['value'], r['ms_per_step'], r['pct_of_8TBps'], r['e2e']['value'])"
