set -u
mkdir -p gpurun_out
T=r02f
timeout 600 python -m pytest tests/test_gpu_prefill.py -m gpu -q 2>&1 | tail -40 > gpurun_out/${T}_pytest_prefill.txt; tail -15 gpurun_out/${T}_pytest_prefill.txt
timeout 600 python -m pytest tests/test_gpu_tune_variants.py tests/test_gpu_backend_multi_query.py -m gpu -q 2>&1 | tail -30 > gpurun_out/${T}_pytest_tune_backend.txt; tail -12 gpurun_out/${T}_pytest_tune_backend.txt
timeout 600 python tools/prefillbench.py --expand --iters 10 > gpurun_out/${T}_prefillbench.jsonl 2> gpurun_out/${T}_prefillbench.err; tail -2 gpurun_out/${T}_prefillbench.err; cat gpurun_out/${T}_prefillbench.jsonl
for t in 0 5 8 9 10 11 12 13 14; do
  timeout 120 python tools/powerprobe.py --config cfg4 --seconds 3 --tune $t >> gpurun_out/${T}_powerprobe_cfg4.jsonl 2>> gpurun_out/${T}_powerprobe.err
  sleep 2
done
for t in 0 1 2 3 4 5 6 7 8 9 10 11 12 13 14; do
  timeout 120 python tools/powerprobe.py --config cfg2 --seconds 2 --tune $t >> gpurun_out/${T}_powerprobe_cfg2.jsonl 2>> gpurun_out/${T}_powerprobe.err
  sleep 1
done
for t in 0 5 10; do
  timeout 120 python tools/powerprobe.py --config cfg3 --seconds 3 --tune $t >> gpurun_out/${T}_powerprobe_cfg3.jsonl 2>> gpurun_out/${T}_powerprobe.err
  sleep 2
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02f_powerprobe_cfg*.jsonl')):
    for l in open(f):
        d=json.loads(l); print(d['config'], d['tune'], d['burst_gbs_10_steps'], d['first_300ms_gbs'], d['last_third_gbs'], d['sm_mhz_last_third'], d['power_w_last_third'])
PY
tail -3 gpurun_out/${T}_powerprobe.err
