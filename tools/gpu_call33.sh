#!/bin/bash
# Head sizes 80 / 112 (16-bit caches) on the stream kernel: parity (sweep + decode_step), timings.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_attention.py tests/test_gpu_decode_step.py -x -q -k "sweep or decode_step" 2>&1 | tail -6 | tee gpurun_out/r01u_pytest_d80_d112.txt
python tools/shapebench.py > gpurun_out/r01u_shapebench.jsonl 2> gpurun_out/r01u_shapebench.err; echo shape=$?
python - <<'PY'
import json
for l in open("gpurun_out/r01u_shapebench.jsonl"):
    r = json.loads(l); print({k: r[k] for k in r if k in ("shape","name","impl","ms","gbs","tbs","max_abs_vs_ref","ms_stream","ms_generic","ms_ref")} or r)
PY
tail -n 3 gpurun_out/r01u_shapebench.err
