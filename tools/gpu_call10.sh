#!/bin/bash
# Evidence refresh for the kernel as committed (in-CTA pre-merge): bench without profiler, then the
# ncu launch list of the same command, then one ncu --set full capture of the main kernel.
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r01b_bench_short.log 2> gpurun_out/r01b_bench_short.err; echo bench=$?
cat gpurun_out/r01b_bench_short.log
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'paged_attention|reshape_and_cache' -c 60 --csv \
  --log-file gpurun_out/r01b_cfg4_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01b_ncu_list.log 2>&1; echo list=$?
tail -9 gpurun_out/r01b_cfg4_launches.csv | awk -F'","' '{print $5, $NF}'
ncu --set full --clock-control none --import-source on -k regex:'paged_attention_stream_kernel' -s 4 -c 1 -f \
  -o gpurun_out/r01b_cfg4_stream_kernel python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01b_ncu_full.log 2>&1; echo full=$?
ls -la gpurun_out/*.ncu-rep
