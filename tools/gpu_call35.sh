#!/bin/bash
# Final evidence for the tree as committed: smoke, bench (both arms), ncu launch list and one
# ncu --set full capture of the stream kernel on the bench command; e5m2 timing row.
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/r01w_smoke.txt
python bench.py --steps 50 --warmup 5 > gpurun_out/r01w_bench_cfg4_n1.json 2> gpurun_out/r01w_bench.err; echo bench=$?
cat gpurun_out/r01w_bench_cfg4_n1.json | cut -c1-1200
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r01w_bench_reference_arm.json 2>> gpurun_out/r01w_bench.err; echo ref=$?
cat gpurun_out/r01w_bench_reference_arm.json | cut -c1-600
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'paged_attention|reshape_and_cache' -c 60 --csv \
  --log-file gpurun_out/r01w_cfg4_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01w_ncu_list.log 2>&1; echo list=$?
tail -4 gpurun_out/r01w_cfg4_launches.csv | awk -F'","' '{print $5, $NF}'
ncu --set full --clock-control none --import-source on -k regex:'paged_attention_stream_kernel' -s 4 -c 1 -f \
  -o gpurun_out/r01w_cfg4_stream_kernel python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01w_ncu_full.log 2>&1; echo full=$?
python tools/shapebench.py --only llama8b_e5m2,llama8b_bs32_fp8 > gpurun_out/r01w_shapebench_e5m2.jsonl 2> gpurun_out/r01w_shapebench.err; echo shape=$?
cut -c1-300 gpurun_out/r01w_shapebench_e5m2.jsonl
tail -n 3 gpurun_out/r01w_bench.err
