#!/bin/bash
# GPU validation: full -m gpu suite with durations, kernel sweep (ours vs the reference's CUDA
# kernels on the same tensors), default bench + reference arm
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider --durations=30 > gpurun_out/pytest_gpu.log 2>&1; echo pytest=$?
tail -45 gpurun_out/pytest_gpu.log
python tools/kbench.py --configs cfg1,cfg2,cfg3,cfg4,cfg5 --ops v1,v2 > gpurun_out/kbench_ours.jsonl 2>gpurun_out/kbench_ours.err; echo kbench=$?
python tools/kbench.py --configs cfg5,cfg4 --tp 8 --ops v2 >> gpurun_out/kbench_ours.jsonl 2>>gpurun_out/kbench_ours.err
cat gpurun_out/kbench_ours.jsonl
timeout 600 python tools/kbench.py --ref-cuda --iters 5 --configs cfg1,cfg2,cfg3,cfg4,cfg5 --ops v1,v2 > gpurun_out/kbench_refcuda.jsonl 2>gpurun_out/kbench_refcuda.err; echo kbench_ref=$?
timeout 300 python tools/kbench.py --ref-cuda --iters 5 --configs cfg5,cfg4 --tp 8 --ops v2 >> gpurun_out/kbench_refcuda.jsonl 2>>gpurun_out/kbench_refcuda.err
cat gpurun_out/kbench_refcuda.jsonl; tail -3 gpurun_out/kbench_refcuda.err
python bench.py > gpurun_out/bench_default.log 2> gpurun_out/bench_default.err; echo bench=$?
cat gpurun_out/bench_default.log
python bench.py --impl reference > gpurun_out/bench_reference.log 2> gpurun_out/bench_reference.err; echo bench_ref=$?
cat gpurun_out/bench_reference.log
