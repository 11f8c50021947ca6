#!/bin/bash
# compute-sanitizer (memcheck, racecheck on shared memory) over the new code paths on small
# shapes; CUDA-graph replay timings of the step next to eager launches.
mkdir -p gpurun_out
SEL='gqa8_fp8_page_edges or gqa4_bf16_page_edges or gqa16_two_tiles_fp8 or d192_f16_fp8 or phi_d96_bf16 or negative_slots or mha_f16_d64_alibi'
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 7 --print-limit 20 \
  python -m pytest tests/test_gpu_decode_step.py tests/test_gpu_flash_layout.py tests/test_gpu_lse.py -x -q -k "$SEL or correct_attn or merge_attn" \
  > gpurun_out/r01l_sanitizer_memcheck.txt 2>&1; echo memcheck=$?
tail -6 gpurun_out/r01l_sanitizer_memcheck.txt
timeout 1500 compute-sanitizer --tool racecheck --error-exitcode 7 --print-limit 20 \
  python -m pytest tests/test_gpu_decode_step.py tests/test_gpu_flash_layout.py -x -q -k "gqa8_fp8_page_edges or gqa4_bf16_page_edges or phi_d96_bf16" \
  > gpurun_out/r01l_sanitizer_racecheck.txt 2>&1; echo racecheck=$?
tail -6 gpurun_out/r01l_sanitizer_racecheck.txt
for spec in "cfg4 8" "cfg5 8" "cfg1 1" "cfg2 1"; do
  set -- $spec
  python tools/kbench.py --configs $1 --tp $2 --ops v2,step,fstep --iters 20 >> gpurun_out/r01l_kbench_eager.jsonl 2>> gpurun_out/r01l_kbench.err
  python tools/kbench.py --configs $1 --tp $2 --ops v2,step,fstep --iters 20 --graph >> gpurun_out/r01l_kbench_graph.jsonl 2>> gpurun_out/r01l_kbench.err
done
cat gpurun_out/r01l_kbench_eager.jsonl gpurun_out/r01l_kbench_graph.jsonl | python -c "
import sys, json
for l in sys.stdin:
    r = json.loads(l); print(r['config'], r['tp'], r['launch'], r['op'], r['ms_median'], r['ms_best'], r.get('gbs'))"
tail -n 5 gpurun_out/r01l_kbench.err
