#!/bin/bash
# fp8 caches at head sizes 80 / 112 on the stream kernel: parity, timings, then the whole GPU suite.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_decode_step.py -x -q -k "sweep or decode_step" 2>&1 | tail -12 | tee gpurun_out/r01y_pytest_fp8_d80_d112.txt
python tools/shapebench.py --only d80_fp8,d112_fp8_gqa4 > gpurun_out/r01y_shapebench.jsonl 2> gpurun_out/r01y_shapebench.err; echo shape=$?
cut -c1-260 gpurun_out/r01y_shapebench.jsonl
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/r01y_pytest_gpu_tail.txt
