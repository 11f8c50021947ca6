#!/bin/bash
# compute-sanitizer passes over smoke-sized cases of every kernel family (memcheck /
# racecheck / synccheck on the small configs).  Run on the GPU box from the repository root:
#   gpurun --timeout 2400 -- 'bash tools/sanitize.sh r03d'
# One log per (tool, target) under gpurun_out/, and TAG_sanitizer_summary.txt with the summary
# lines; copy the summary (and any log with findings) into profiles/.
set -u
TAG=${1:-san}
mkdir -p gpurun_out
SUM=gpurun_out/${TAG}_sanitizer_summary.txt
: > $SUM
CS=/usr/local/cuda/bin/compute-sanitizer
run() {  # run TOOL NAME LIMIT cmd...
  local tool=$1 name=$2 limit=$3; shift 3
  local log=gpurun_out/${TAG}_${tool}_${name}.txt
  timeout $limit $CS --tool $tool --print-limit 20 --log-file $log "$@" > gpurun_out/${TAG}_${tool}_${name}.out 2>&1
  local rc=$?
  {
    echo "== $tool / $name: exit $rc ($(tail -1 gpurun_out/${TAG}_${tool}_${name}.out | cut -c1-120))"
    grep -E "ERROR SUMMARY|RACECHECK SUMMARY|Internal Sanitizer Error|hazard|Invalid|Barrier error" $log | sort | uniq -c | head -12
  } >> $SUM
}
SMOKE=(python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')")
# ragged + empty sequences, cut sequences (tickets), fused write, flash layout, chunk kernel, the
# built-in scratch entry point: a few small cases per family
CASES=(python -m pytest -x -q -m gpu -p no:cacheprovider
       "tests/test_gpu_attention.py::test_zero_length_sequence_v1_gives_zeros"
       "tests/test_gpu_attention.py::test_c_abi_without_workspace_uses_generic_and_stream_needs_one"
       "tests/test_gpu_prefill.py::test_long_chunk_with_context_and_empty_sequences"
       "tests/test_gpu_prefill.py::test_sequence_shorter_than_its_chunk_gets_a_defined_result"
       "tests/test_gpu_lse.py::test_lse_call_leaves_plain_v1_result_unchanged"
       "tests/test_gpu_lse.py::test_correct_attn_out_many_ranks_and_empty_rows"
       "tests/test_gpu_decode_step.py::test_decode_step_skips_negative_slots"
       "tests/test_gpu_decode_step.py::test_decode_step_lse_and_strided_key_value"
       "tests/test_gpu_flash_layout.py::test_flash_layout_backend_decode_step_matches_oracle"
       "tests/test_zz_no_alloc.py")
run memcheck smoke 900 "${SMOKE[@]}"
run memcheck cases 1200 "${CASES[@]}"
run racecheck smoke 1200 "${SMOKE[@]}"
run synccheck smoke 900 "${SMOKE[@]}"
run racecheck cases 1500 "${CASES[@]}"
cat $SUM
