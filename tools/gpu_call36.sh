#!/bin/bash
# Same-box scaling series on the final tree, launched as the driver launches it: N = 1, 2, 4, 8.
mkdir -p gpurun_out
python bench.py --gpus 1 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r01x_scale_n1.json 2> gpurun_out/r01x_scale_n1.err; echo n1=$?
for N in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2961$N \
    bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/r01x_scale_n$N.json 2> gpurun_out/r01x_scale_n$N.err; echo n$N=$?
done
python - <<'PY'
import json
for n in ("n1", "n2", "n4", "n8"):
    try:
        r = json.loads(open(f"gpurun_out/r01x_scale_{n}.json").read().strip().splitlines()[-1])
        print(n, r["value"], r["ms_per_step"], r["roofline"]["kernel_ms"], r["e2e"]["value"], (r["e2e"]["graph_replay"] or {}).get("value"), r["clocks"])
    except Exception as e:
        print(n, "failed", e)
PY
grep -v "OMP_NUM_THREADS\|^\*\*\*" gpurun_out/r01x_scale_n8.err | tail -n 3
