#!/bin/bash
# Full GPU suite + evidence for the kernel as committed (fused decode_step is the bench step):
# bench without profiler, ncu launch list of the same command, one ncu --set full capture of the
# stream kernel, reference arm.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/r01g_pytest_gpu_tail.txt
python bench.py --steps 50 --warmup 5 > gpurun_out/r01g_bench_cfg4_n1.json 2> gpurun_out/r01g_bench.err; echo bench=$?
cat gpurun_out/r01g_bench_cfg4_n1.json
python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r01g_bench_short.json 2>> gpurun_out/r01g_bench.err; echo short=$?
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'paged_attention|reshape_and_cache' -c 60 --csv \
  --log-file gpurun_out/r01g_cfg4_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01g_ncu_list.log 2>&1; echo list=$?
tail -9 gpurun_out/r01g_cfg4_launches.csv | awk -F'","' '{print $5, $NF}'
ncu --set full --clock-control none --import-source on -k regex:'paged_attention_stream_kernel' -s 4 -c 1 -f \
  -o gpurun_out/r01g_cfg4_stream_kernel python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01g_ncu_full.log 2>&1; echo full=$?
ls -la gpurun_out/*.ncu-rep
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r01g_bench_reference_arm.json 2>> gpurun_out/r01g_bench.err; echo ref=$?
cat gpurun_out/r01g_bench_reference_arm.json
tail -n 5 gpurun_out/r01g_bench.err
