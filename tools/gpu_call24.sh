#!/bin/bash
# Config 5 as BASELINE.json states it (ragged 128..131072, permuted block tables, 8 x B200) and
# config 4 at N=8, launched as the driver launches bench.py.
mkdir -p gpurun_out
nvidia-smi -L | wc -l
for C in cfg5 cfg4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 \
    bench.py --gpus 8 --config $C --steps 50 --warmup 5 > gpurun_out/r01p_bench_${C}_n8.json 2> gpurun_out/r01p_bench_${C}_n8.err; echo $C=$?
done
python - <<'PY'
import json
for c in ("cfg5", "cfg4"):
    try:
        r = json.loads(open(f"gpurun_out/r01p_bench_{c}_n8.json").read().strip().splitlines()[-1])
        print(c, r["value"], r["ms_per_step"], r["roofline"]["achieved"], r["roofline"]["kernel_ms"], r["e2e"]["value"], r["clocks"])
    except Exception as e:
        print(c, "failed", e)
PY
tail -n 3 gpurun_out/r01p_bench_cfg5_n8.err
