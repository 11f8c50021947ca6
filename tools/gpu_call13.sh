#!/bin/bash
# flash-layout decode kernel: parity, then timings vs the x-packed kernel; correct_attn_out v3.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_flash_layout.py tests/test_gpu_lse.py tests/test_gpu_decode_step.py -x -q 2>&1 | tail -25 | tee gpurun_out/r01e_pytest_flash.txt
for spec in "cfg4 8" "cfg2 1" "cfg3 1" "cfg4 1"; do
  set -- $spec
  timeout 300 python tools/kbench.py --configs $1 --tp $2 --ops v2,flash_nhd,flash_hnd,fflash_nhd,v2,flash_nhd --iters 20 >> gpurun_out/r01e_kbench_flash.jsonl 2>> gpurun_out/r01e_kbench.err
done
cat gpurun_out/r01e_kbench_flash.jsonl | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l); print(r['config'], r['tp'], r['op'], r['ms_median'], r['ms_best'], r.get('gbs'))"
python tools/opbench.py --iters 20 > gpurun_out/r01e_opbench_after.jsonl 2> gpurun_out/r01e_opbench.err
grep -h "merge_attn_states\|correct_attn_out" gpurun_out/r01e_opbench_after.jsonl | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l); print(r['case'], r['impl'], r['us_median'], r['gbs'], r['frac_measured_hbm'])"
tail -n 5 gpurun_out/r01e_kbench.err gpurun_out/r01e_opbench.err
