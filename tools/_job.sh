set -u
mkdir -p gpurun_out
T=r02n
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -8 > gpurun_out/${T}_pytest.txt; tail -4 gpurun_out/${T}_pytest.txt | cut -c1-200
tools/gpu_job.sh $T smoke | tee gpurun_out/${T}_smoke.txt
python bench.py --impl reference > gpurun_out/${T}_bench_reference_arm.json 2> gpurun_out/${T}_bench_ref.err; cut -c1-300 gpurun_out/${T}_bench_reference_arm.json
python bench.py > gpurun_out/${T}_bench_cfg4.json 2> gpurun_out/${T}_bench_cfg4.err; echo "bench rc=$?"; tail -2 gpurun_out/${T}_bench_cfg4.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02n_bench_cfg4.json'))
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],'sustained',d['sustained']['gbs'],'frac',d['roofline']['frac'],'parity',d['parity']['ok'], 'launches', d['gpu_launches'])
print({k:(v['gbs'],v['ms_per_step'],v['parity'].get('ok')) for k,v in d['configs'].items()})
PY
tools/gpu_job.sh $T list cfg4
cap() {  # name, kernel regex, skip, command...
  local name=$1 rx=$2 skip=$3; shift 3
  "$@" > gpurun_out/${T}_${name}_plain.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:"$rx" -s $skip -c 1 -f \
      -o gpurun_out/${T}_${name} "$@" > gpurun_out/${T}_${name}_ncu.log 2>&1
  echo "cap $name rc=$?"
  if [ -f gpurun_out/${T}_${name}.ncu-rep ]; then
    ncu -i gpurun_out/${T}_${name}.ncu-rep --page raw --csv > gpurun_out/${T}_${name}_raw.csv 2>/dev/null
    python tools/ncu_summary.py gpurun_out/${T}_${name}_raw.csv > gpurun_out/${T}_${name}_summary.txt 2>&1
    rm -f gpurun_out/${T}_${name}.ncu-rep
  fi
  rm -f gpurun_out/${T}_${name}_plain.log
}
cap cfg4_block paged_attention_nhd_kernel 4 python bench.py --config cfg4 --steps 3 --warmup 3 --no-cpu-baseline --no-check --sustain 0 --no-side-configs
cap cfg3_nhd paged_attention_nhd_kernel 3 python tools/kbench.py --configs cfg3 --ops flash_nhd --iters 3
grep -E "dram__bytes_read.sum \[|gpu__time_duration|dram__cycles_active.avg.pct|issue_active.avg.pct|registers_per_thread \[|tensor_cycles_active.avg.pct_of_peak_sustained_active" gpurun_out/${T}_cfg4_block_summary.txt gpurun_out/${T}_cfg3_nhd_summary.txt | cut -c1-200
