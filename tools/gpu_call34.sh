#!/bin/bash
# fp8_e5m2 on the stream kernel: parity tests (attention + fused decode_step), then the whole GPU suite.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_decode_step.py -x -q -k "e5m2 or other_dtypes" 2>&1 | tail -15 | tee gpurun_out/r01v_pytest_e5m2.txt
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/r01v_pytest_gpu_tail.txt
