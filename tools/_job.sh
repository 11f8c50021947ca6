set -u
mkdir -p gpurun_out
T=r02k
timeout 600 python -m pytest tests/test_gpu_flash_layout.py -m gpu -q -x 2>&1 | tail -30 > gpurun_out/${T}_pytest_flash.txt; tail -30 gpurun_out/${T}_pytest_flash.txt | cut -c1-220
for t in 0 99; do
  B200PA_STREAM_TUNE=$t timeout 300 python tools/kbench.py --configs cfg3,cfg2,cfg4 --ops flash_nhd,flash_hnd,v2 --iters 10 >> gpurun_out/${T}_kbench_nhd.jsonl 2>> gpurun_out/${T}_kbench.err
done
cut -c1-260 gpurun_out/${T}_kbench_nhd.jsonl; tail -3 gpurun_out/${T}_kbench.err
