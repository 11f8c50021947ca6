#!/bin/bash
# GPU call: kernel sweep over the 5 configs, then ncu launch list + full capture on cfg4.
mkdir -p gpurun_out
python -m pytest tests/test_gpu_attention.py -m gpu -q -k "cuda_graph" -p no:cacheprovider > gpurun_out/pytest_graph.log 2>&1; echo graph_test=$?
python tools/kbench.py --iters 20 > gpurun_out/kbench.log 2> gpurun_out/kbench.err; echo kbench=$?
cat gpurun_out/kbench.log
python tools/kbench.py --iters 20 --tp 8 --configs cfg4,cfg5 > gpurun_out/kbench_tp8.log 2>> gpurun_out/kbench.err; echo kbench_tp8=$?
cat gpurun_out/kbench_tp8.log
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'paged_attention|reshape_and_cache' -c 60 \
    --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo ncu_launch=$?
ncu --set full --clock-control none --import-source on -k regex:paged_attention_fast -s 4 -c 2 \
    -o gpurun_out/prof_fast_cfg4 $CMD > gpurun_out/ncu_full.log 2>&1
echo ncu_full=$?
tail -3 gpurun_out/ncu_full.log
