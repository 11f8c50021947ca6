#!/bin/bash
# stream kernel on 32-token cache blocks: parity (sweep + decode_step) and timings.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_attention.py tests/test_gpu_decode_step.py tests/test_gpu_flash_layout.py tests/test_gpu_lse.py -x -q 2>&1 | tail -15 | tee gpurun_out/r01k_pytest_bs32.txt
timeout 600 python tools/shapebench.py > gpurun_out/r01k_shapebench.jsonl 2> gpurun_out/r01k_shapebench.err
cat gpurun_out/r01k_shapebench.jsonl | cut -c1-260; tail -n 5 gpurun_out/r01k_shapebench.err
