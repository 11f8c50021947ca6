#!/bin/bash
# stream kernel on head sizes 32 / 96 / 192 / 256: parity tests + timings vs generic and reference CUDA.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_attention.py tests/test_gpu_decode_step.py -x -q 2>&1 | tail -15 | tee gpurun_out/r01j_pytest_head_sizes.txt
timeout 600 python tools/shapebench.py > gpurun_out/r01j_shapebench.jsonl 2> gpurun_out/r01j_shapebench.err
cat gpurun_out/r01j_shapebench.jsonl; tail -n 5 gpurun_out/r01j_shapebench.err
