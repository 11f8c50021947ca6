#!/bin/bash
# flash-layout kernel with cp.async granules instead of per-row TMA copies: parity + timing.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_flash_layout.py -x -q 2>&1 | tail -25 | tee gpurun_out/r01f_pytest_flash.txt
for spec in "cfg4 8" "cfg2 1" "cfg3 1" "cfg4 1" "cfg5 8" "cfg1 1"; do
  set -- $spec
  timeout 300 python tools/kbench.py --configs $1 --tp $2 --ops v2,flash_nhd,flash_hnd,fflash_nhd,v2,flash_nhd --iters 20 >> gpurun_out/r01f_kbench_flash.jsonl 2>> gpurun_out/r01f_kbench.err
done
cat gpurun_out/r01f_kbench_flash.jsonl | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l); print(r['config'], r['tp'], r['op'], r['ms_median'], r['ms_best'], r.get('gbs'))"
tail -n 5 gpurun_out/r01f_kbench.err
