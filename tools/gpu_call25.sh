#!/bin/bash
# ncu --set full (with source) of the stream kernel and its merge kernel on the cfg5 tp8 shard
# (ragged, 0.4 GB of KV: the regime where the step is furthest from the HBM roofline).
mkdir -p gpurun_out
python tools/kbench.py --configs cfg5 --tp 8 --ops v2 --iters 20 > gpurun_out/r01q_kbench_cfg5_tp8.jsonl 2> gpurun_out/r01q_kbench.err; echo kbench=$?
cat gpurun_out/r01q_kbench_cfg5_tp8.jsonl | cut -c1-300
ncu --set full --clock-control none --import-source on -k regex:'paged_attention_stream' -s 8 -c 2 -f \
  -o gpurun_out/r01q_cfg5_tp8_stream python tools/kbench.py --configs cfg5 --tp 8 --ops v2 --iters 3 > gpurun_out/r01q_ncu_full.log 2>&1; echo full=$?
ls -la gpurun_out/*.ncu-rep
tail -n 3 gpurun_out/r01q_ncu_full.log
