#!/bin/bash
# Same-box scaling series exactly as the driver launches it: N = 1, 2, 4, 8 back to back on one
# 8-GPU box (fused decode_step is the step), then N = 8 with the two-op step for the A/B.
mkdir -p gpurun_out
nvidia-smi -L | head -8
python bench.py --gpus 1 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r01h_scale_n1.json 2> gpurun_out/r01h_scale_n1.err; echo n1=$?
for N in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N \
    bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/r01h_scale_n$N.json 2> gpurun_out/r01h_scale_n$N.err; echo n$N=$?
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 \
  bench.py --gpus 8 --steps 50 --warmup 5 --unfused > gpurun_out/r01h_scale_n8_unfused.json 2> gpurun_out/r01h_scale_n8_unfused.err; echo n8u=$?
python bench.py --gpus 1 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r01h_scale_n1_again.json 2>> gpurun_out/r01h_scale_n1.err; echo n1b=$?
python - <<'PY'
import json
for n in ("n1", "n2", "n4", "n8", "n8_unfused", "n1_again"):
    try:
        r = json.loads(open(f"gpurun_out/r01h_scale_{n}.json").read().strip().splitlines()[-1])
        print(n, r["value"], r["ms_per_step"], r["roofline"]["kernel_ms"], r["e2e"]["value"], r["clocks"])
    except Exception as e:
        print(n, "failed", e)
PY
tail -n 3 gpurun_out/r01h_scale_n8.err
