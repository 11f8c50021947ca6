#!/bin/bash
# decode_step (fused cache write) first GPU check: parity tests, then same-process op timings
# step (two ops) vs fstep (one launch), the LSE merge kernels after their rewrite, short bench A/B.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_decode_step.py tests/test_gpu_lse.py -x -q 2>&1 | tail -15 | tee gpurun_out/r01c_pytest_decode_step.txt
for spec in "cfg4 8" "cfg4 1" "cfg1 1" "cfg5 8" "cfg3 1"; do
  set -- $spec
  python tools/kbench.py --configs $1 --tp $2 --ops v2,step,fstep,step,fstep --iters 20 >> gpurun_out/r01c_kbench_fstep.jsonl 2>> gpurun_out/r01c_kbench.err
done
cat gpurun_out/r01c_kbench_fstep.jsonl | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l); print(r['config'], r['tp'], r['op'], r['ms_median'], r['ms_best'], r.get('gbs'))"
python tools/opbench.py --iters 20 > gpurun_out/r01c_opbench_after.jsonl 2> gpurun_out/r01c_opbench.err
grep -h "merge_attn_states\|correct_attn_out" gpurun_out/r01c_opbench_after.jsonl | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l); print(r['case'], r['impl'], r['us_median'], r['gbs'], r['frac_measured_hbm'])"
python bench.py --steps 30 --warmup 3 --no-cpu-baseline > gpurun_out/r01c_bench_fused.json 2> gpurun_out/r01c_bench_fused.err; echo fused=$?
python bench.py --steps 30 --warmup 3 --no-cpu-baseline --unfused > gpurun_out/r01c_bench_unfused.json 2> gpurun_out/r01c_bench_unfused.err; echo unfused=$?
python -c "
import json
for n in ('fused','unfused'):
    r = json.load(open(f'gpurun_out/r01c_bench_{n}.json')); print(n, r['value'], r['ms_per_step'], r['roofline']['kernel_ms'], r['e2e']['value'], r['gpu_launches'])"
tail -3 gpurun_out/r01c_kbench.err gpurun_out/r01c_opbench.err gpurun_out/r01c_bench_fused.err
