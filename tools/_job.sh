set -u
mkdir -p gpurun_out
T=r02g
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/${T}_pytest.txt; cat gpurun_out/${T}_pytest.txt
tools/gpu_job.sh $T smoke
python bench.py > gpurun_out/${T}_bench_cfg4.json 2> gpurun_out/${T}_bench_cfg4.err; echo "bench rc=$?"; tail -3 gpurun_out/${T}_bench_cfg4.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02g_bench_cfg4.json'))
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],'sustained',d['sustained']['gbs'],'frac',d['roofline']['frac'],'parity',d['parity']['ok'])
print({k:(v['gbs'],v['ms_per_step'],v['parity'].get('ok')) for k,v in d['configs'].items()})
PY
python tools/opbench.py > gpurun_out/${T}_opbench.jsonl 2> gpurun_out/${T}_opbench.err; tail -2 gpurun_out/${T}_opbench.err; grep reshape_and_cache gpurun_out/${T}_opbench.jsonl | cut -c1-260
for t in 0 5 8 9 11; do
  B200PA_STREAM_TUNE=$t python tools/kbench.py --configs cfg5,cfg4 --tp 8 --ops fstep --iters 30 >> gpurun_out/${T}_kbench_small.jsonl 2>> gpurun_out/${T}_kbench.err
done
for t in 0 1 8 9 10 12; do
  B200PA_STREAM_TUNE=$t python tools/kbench.py --configs cfg1,cfg2 --tp 1 --ops fstep --iters 30 >> gpurun_out/${T}_kbench_small.jsonl 2>> gpurun_out/${T}_kbench.err
done
cut -c1-230 gpurun_out/${T}_kbench_small.jsonl; tail -2 gpurun_out/${T}_kbench.err
# ---- ncu --set full captures (each only after the same command exited 0 without ncu)
cap() {  # name, kernel regex, skip, command...
  local name=$1 rx=$2 skip=$3; shift 3
  "$@" > gpurun_out/${T}_${name}_plain.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:"$rx" -s $skip -c 1 -f \
      -o gpurun_out/${T}_${name} "$@" > gpurun_out/${T}_${name}_ncu.log 2>&1
  echo "cap $name rc=$?"
}
cap cfg4_stream paged_attention_stream_kernel 4 python bench.py --config cfg4 --steps 3 --warmup 3 --no-cpu-baseline --no-check --sustain 0 --no-side-configs
cap cfg2_stream paged_attention_stream_kernel 4 python bench.py --config cfg2 --steps 3 --warmup 3 --no-cpu-baseline --no-check --sustain 0 --no-side-configs
cap d192_fp8 paged_attention_stream_kernel 3 python tools/shapebench.py --only d192_fp8 --iters 2
cap d64_fp8_gqa8 paged_attention_stream_kernel 3 python tools/shapebench.py --only d64_fp8_gqa8 --iters 2
cap d112_fp8_gqa4 paged_attention_stream_kernel 3 python tools/shapebench.py --only d112_fp8_gqa4 --iters 2
cap flash_nhd paged_attention_stream_kernel 3 python tools/kbench.py --configs cfg3 --ops flash_nhd --iters 3
cap prefill_fp8 paged_prefill_kernel 2 python tools/prefillbench.py --cases chunk2k_ctx6k_fp8 --iters 2
cap prefill_bf16 paged_prefill_kernel 2 python tools/prefillbench.py --cases chunk2k_ctx6k_bf16 --iters 2
ls -la gpurun_out/*.ncu-rep | awk '{print $5, $9}'
