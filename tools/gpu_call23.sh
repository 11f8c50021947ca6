#!/bin/bash
# bench.py lines for the other BASELINE configs on 1 GPU (cfg1, cfg2, cfg3, cfg5): same contract
# as the default cfg4 line, whole step = fused decode_step, no CPU leg.
mkdir -p gpurun_out
for C in cfg1 cfg2 cfg3 cfg5; do
  python bench.py --config $C --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r01o_bench_${C}_n1.json 2> gpurun_out/r01o_bench_${C}.err; echo $C=$?
done
python - <<'PY'
import json
for c in ("cfg1", "cfg2", "cfg3", "cfg5"):
    try:
        r = json.loads(open(f"gpurun_out/r01o_bench_{c}_n1.json").read().strip().splitlines()[-1])
        print(c, r["value"], r["ms_per_step"], r["roofline"]["achieved"], r["roofline"]["kernel_ms"], r["e2e"]["value"], r["clocks"])
    except Exception as e:
        print(c, "failed", e)
PY
tail -n 3 gpurun_out/r01o_bench_cfg5.err
