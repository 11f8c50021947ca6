#!/bin/bash
# Tuning-variant sweep (B200PA_STREAM_TUNE) on the small shards: cfg1 (bf16), cfg5 tp8 and cfg4 tp8 (fp8).
mkdir -p gpurun_out
: > gpurun_out/r01t_kbench_tune_small.jsonl
for spec in "cfg1 1" "cfg5 8" "cfg4 8"; do
  set -- $spec
  for T in 0 1 2 3 4 5 6 7 0; do
    B200PA_STREAM_TUNE=$T python tools/kbench.py --configs $1 --tp $2 --ops v2,fstep --iters 30 >> gpurun_out/r01t_kbench_tune_small.jsonl 2>> gpurun_out/r01t_kbench.err
  done
done
python - <<'PY'
import json
for l in open("gpurun_out/r01t_kbench_tune_small.jsonl"):
    r = json.loads(l); print(r["config"], r["tp"], r["tune"], r["op"], r["ms_median"], r["ms_best"], r.get("gbs"))
PY
tail -n 3 gpurun_out/r01t_kbench.err
