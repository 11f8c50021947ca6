#!/bin/bash
# ncu --set full of the merge kernel on cfg4's tp=8 shard (the shard the N=8 scaling run times)
mkdir -p gpurun_out
python tools/kbench.py --configs cfg4 --tp 8 --ops v2 --iters 5 > gpurun_out/merge_plain.log 2>&1; echo plain=$?
ncu --set full --clock-control none --import-source on -k regex:'stream_merge' -s 4 -c 1 -f -o gpurun_out/merge_cfg4_tp8 \
  python tools/kbench.py --configs cfg4 --tp 8 --ops v2 --iters 5 > gpurun_out/merge_ncu.log 2>&1; echo ncu=$?
ls -la gpurun_out/*.ncu-rep
