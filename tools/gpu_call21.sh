#!/bin/bash
# smoke() with the fused and flash paths; bench with the single-copy e2e staging buffer.
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
python bench.py --steps 50 --warmup 5 > gpurun_out/r01m_bench_cfg4_n1.json 2> gpurun_out/r01m_bench.err; echo bench=$?
python -c "
import json
r = json.load(open('gpurun_out/r01m_bench_cfg4_n1.json')); print(r['value'], r['ms_per_step'], r['roofline']['kernel_ms'], r['e2e'], r['gpu_launches'])"
tail -n 3 gpurun_out/r01m_bench.err
timeout 900 python -m pytest tests/test_gpu_backend_multi_query.py tests/test_gpu_vs_reference_cuda.py tests/test_gpu_block_ops.py tests/test_gpu_cache.py -x -q 2>&1 | tail -4
