#!/bin/bash
# multi-query backend test; generic (SIMT, whole-domain) kernel vs the reference's CUDA kernels.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_backend_multi_query.py -x -q 2>&1 | tail -15 | tee gpurun_out/r01i_pytest_multi_query.txt
timeout 600 python tools/kbench.py --impl 1 --configs cfg1,cfg2,cfg3 --ops v1,v2 --iters 10 > gpurun_out/r01i_kbench_generic.jsonl 2> gpurun_out/r01i_kbench.err
timeout 600 python tools/kbench.py --ref-cuda --configs cfg1,cfg2,cfg3 --ops v1,v2 --iters 10 > gpurun_out/r01i_kbench_refcuda.jsonl 2>> gpurun_out/r01i_kbench.err
cat gpurun_out/r01i_kbench_generic.jsonl gpurun_out/r01i_kbench_refcuda.jsonl | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l); print(r['config'], r['impl'], r['op'], r['ms_median'], r.get('gbs'))"
tail -n 5 gpurun_out/r01i_kbench.err
