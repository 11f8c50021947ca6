set -u
mkdir -p gpurun_out
T=r02h
python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/${T}_pytest.txt; tail -25 gpurun_out/${T}_pytest.txt
python tools/opbench.py 2> gpurun_out/${T}_opbench.err | grep reshape_and_cache > gpurun_out/${T}_opbench_cache.jsonl; grep '"reshape_and_cache ' gpurun_out/${T}_opbench_cache.jsonl | grep 16384 | cut -c1-200
# ---- ncu --set full captures (each only after the same command exited 0 without ncu); the
# reports are exported to CSV on the box (gpurun_out is capped at 64 MiB)
cap() {  # name, kernel regex, skip, keep-report(0/1), command...
  local name=$1 rx=$2 skip=$3 keep=$4; shift 4
  "$@" > gpurun_out/${T}_${name}_plain.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:"$rx" -s $skip -c 1 -f \
      -o gpurun_out/${T}_${name} "$@" > gpurun_out/${T}_${name}_ncu.log 2>&1
  echo "cap $name rc=$?"
  if [ -f gpurun_out/${T}_${name}.ncu-rep ]; then
    ncu -i gpurun_out/${T}_${name}.ncu-rep --page raw --csv > gpurun_out/${T}_${name}_raw.csv 2>/dev/null
    python tools/ncu_summary.py gpurun_out/${T}_${name}_raw.csv > gpurun_out/${T}_${name}_summary.txt 2>&1
    if [ $keep = 0 ]; then rm -f gpurun_out/${T}_${name}.ncu-rep; fi
  fi
  rm -f gpurun_out/${T}_${name}_plain.log
}
cap cfg4_stream paged_attention_stream_kernel 4 1 python bench.py --config cfg4 --steps 3 --warmup 3 --no-cpu-baseline --no-check --sustain 0 --no-side-configs
cap cfg2_stream paged_attention_stream_kernel 4 0 python bench.py --config cfg2 --steps 3 --warmup 3 --no-cpu-baseline --no-check --sustain 0 --no-side-configs
cap d192_fp8 paged_attention_stream_kernel 3 0 python tools/shapebench.py --only d192_fp8 --iters 2
cap d64_fp8_gqa8 paged_attention_stream_kernel 3 0 python tools/shapebench.py --only d64_fp8_gqa8 --iters 2
cap d112_fp8_gqa4 paged_attention_stream_kernel 3 0 python tools/shapebench.py --only d112_fp8_gqa4 --iters 2
cap flash_nhd paged_attention_stream_kernel 3 0 python tools/kbench.py --configs cfg3 --ops flash_nhd --iters 3
cap prefill_fp8 paged_prefill_kernel 2 1 python tools/prefillbench.py --cases chunk2k_ctx6k_fp8 --iters 2
cap prefill_bf16 paged_prefill_kernel 2 0 python tools/prefillbench.py --cases chunk2k_ctx6k_bf16 --iters 2
cap cache_run reshape_and_cache_run_kernel 3 0 python tools/opbench.py --iters 3
du -sh gpurun_out; ls gpurun_out | head -50
