#!/bin/bash
# Full GPU suite on the tree as committed (flash head sizes, four translation units, smoke with
# decode_step + flash) + bench + ncu launch list of the same command.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/r01n_pytest_gpu_tail.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/r01n_smoke.txt
python bench.py --steps 50 --warmup 5 > gpurun_out/r01n_bench_cfg4_n1.json 2> gpurun_out/r01n_bench.err; echo bench=$?
cat gpurun_out/r01n_bench_cfg4_n1.json
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'paged_attention|reshape_and_cache' -c 60 --csv \
  --log-file gpurun_out/r01n_cfg4_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01n_ncu_list.log 2>&1; echo list=$?
tail -6 gpurun_out/r01n_cfg4_launches.csv | awk -F'","' '{print $5, $NF}'
tail -n 5 gpurun_out/r01n_bench.err
