#!/bin/bash
# Two GPUs in one box: the two-devices-in-one-process test (per-device shared-memory opt-in of the
# stream kernel), then the driver's N=2 launch line with the graph-replay e2e leg.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_attention.py -x -q -k "second_device" 2>&1 | tail -3 | tee gpurun_out/r01r_pytest_second_device.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 \
  bench.py --gpus 2 --steps 50 --warmup 5 > gpurun_out/r01r_bench_cfg4_n2.json 2> gpurun_out/r01r_bench_cfg4_n2.err; echo n2=$?
python bench.py --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r01r_bench_cfg4_n1.json 2> gpurun_out/r01r_bench_cfg4_n1.err; echo n1=$?
python - <<'PY'
import json
for n in ("n2", "n1"):
    try:
        r = json.loads(open(f"gpurun_out/r01r_bench_cfg4_{n}.json").read().strip().splitlines()[-1])
        print(n, r["value"], r["ms_per_step"], r["roofline"]["kernel_ms"], r["e2e"]["value"], r["e2e"]["ms_per_step"], r["e2e"]["graph_replay"])
    except Exception as e:
        print(n, "failed", e)
PY
grep -v "OMP_NUM_THREADS\|^\*\*\*" gpurun_out/r01r_bench_cfg4_n2.err | tail -n 4
tail -n 3 gpurun_out/r01r_bench_cfg4_n1.err
