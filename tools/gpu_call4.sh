#!/bin/bash
# full GPU validation of the stream kernel + bench + ncu evidence (one ncu tool per call)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo pytest=$?
tail -4 gpurun_out/pytest_gpu.log
python bench.py --steps 50 --warmup 5 > gpurun_out/bench_default.log 2> gpurun_out/bench_default.err; echo bench=$?
cat gpurun_out/bench_default.log
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'paged_attention|reshape_and_cache' -c 60 \
    --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo ncu_launch=$?
ncu --set full --clock-control none --import-source on -k regex:paged_attention_stream_kernel -s 4 -c 2 \
    -o gpurun_out/prof_stream_cfg4 $CMD > gpurun_out/ncu_full.log 2>&1
echo ncu_full=$?
tail -3 gpurun_out/ncu_full.log
