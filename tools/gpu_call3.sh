#!/bin/bash
# dev GPU call: stream kernel correctness vs the generic kernel, then timing sweep
mkdir -p gpurun_out
timeout 600 python tools/dev_stream_check.py --impls 3 > gpurun_out/stream_check.log 2>&1; echo check=$?
cat gpurun_out/stream_check.log | tail -20
timeout 120 ./tools/bwprobe > gpurun_out/bwprobe.log 2>&1; cat gpurun_out/bwprobe.log
: > gpurun_out/kbench_stream.log
for t in 0 4 6 7; do
  B200PA_STREAM_TUNE=$t timeout 300 python tools/kbench.py --iters 20 --ops v2 --configs cfg2,cfg3,cfg4 >> gpurun_out/kbench_stream.log 2>gpurun_out/kbench_stream.err
  B200PA_STREAM_TUNE=$t timeout 300 python tools/kbench.py --iters 20 --ops v2 --tp 8 --configs cfg4,cfg5 >> gpurun_out/kbench_stream.log 2>>gpurun_out/kbench_stream.err
done
timeout 300 python tools/kbench.py --iters 20 --ops v1,v2 --configs cfg1,cfg2,cfg5 >> gpurun_out/kbench_stream.log 2>>gpurun_out/kbench_stream.err
cat gpurun_out/kbench_stream.log
tail -3 gpurun_out/kbench_stream.err
