#!/bin/bash
# GPU validation after the LSE row: new LSE tests first (fast feedback), then the whole -m gpu
# suite, then the default bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_lse.py -q -p no:cacheprovider -x > gpurun_out/pytest_lse.log 2>&1; echo pytest_lse=$?
tail -30 gpurun_out/pytest_lse.log
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider --durations=15 > gpurun_out/pytest_gpu.log 2>&1; echo pytest=$?
tail -30 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_default.log 2> gpurun_out/bench_default.err; echo bench=$?
cat gpurun_out/bench_default.log
