#!/bin/bash
# decode_step v2 (new token prefetched into registers at begin_piece) + correct_attn_out with
# batched LSE loads: parity tests, op timings step vs fstep, opbench LSE rows.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_decode_step.py tests/test_gpu_lse.py -x -q 2>&1 | tail -15 | tee gpurun_out/r01d_pytest_decode_step.txt
for spec in "cfg4 8" "cfg5 8" "cfg1 1" "cfg2 1" "cfg3 1" "cfg5 1" "cfg4 1"; do
  set -- $spec
  python tools/kbench.py --configs $1 --tp $2 --ops v2,step,fstep,step,fstep,v2 --iters 20 >> gpurun_out/r01d_kbench_fstep.jsonl 2>> gpurun_out/r01d_kbench.err
done
cat gpurun_out/r01d_kbench_fstep.jsonl | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l); print(r['config'], r['tp'], r['op'], r['ms_median'], r['ms_best'], r.get('gbs'))"
python tools/opbench.py --iters 20 > gpurun_out/r01d_opbench_after.jsonl 2> gpurun_out/r01d_opbench.err
grep -h "merge_attn_states\|correct_attn_out" gpurun_out/r01d_opbench_after.jsonl | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l); print(r['case'], r['impl'], r['us_median'], r['gbs'], r['frac_measured_hbm'])"
tail -n 3 gpurun_out/r01d_kbench.err gpurun_out/r01d_opbench.err
