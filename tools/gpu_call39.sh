#!/bin/bash
# Last call of round 1: whole GPU suite, smoke and the default bench line on the tree as committed.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/r01end_pytest_gpu_tail.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1 | tee gpurun_out/r01end_smoke.txt
python bench.py > gpurun_out/r01end_bench_cfg4_n1.json 2> gpurun_out/r01end_bench.err; echo bench=$?
python -c "
import json; r=json.load(open('gpurun_out/r01end_bench_cfg4_n1.json')); print(r['value'], r['ms_per_step'], r['roofline']['achieved'], r['roofline']['frac'], r['e2e']['value'], r['e2e']['graph_replay'], r['cpu_baseline']['value'], r['gpu_launches'], r['clocks'])"
tail -n 3 gpurun_out/r01end_bench.err
