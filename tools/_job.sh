set -u
mkdir -p gpurun_out
T=r02e
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/${T}_pytest.txt; cat gpurun_out/${T}_pytest.txt
tools/gpu_job.sh $T smoke
python bench.py > gpurun_out/${T}_bench_cfg4.json 2> gpurun_out/${T}_bench_cfg4.err; echo "bench rc=$?"; tail -3 gpurun_out/${T}_bench_cfg4.err; cut -c1-3000 gpurun_out/${T}_bench_cfg4.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_reference_arm.json 2>> gpurun_out/${T}_bench_cfg4.err; cut -c1-600 gpurun_out/${T}_bench_reference_arm.json
for t in 0 1 2 3 4 5 6 7; do
  python tools/powerprobe.py --config cfg4 --seconds 3 --tune $t >> gpurun_out/${T}_powerprobe_tunes.jsonl 2>> gpurun_out/${T}_powerprobe.err
  sleep 2
done
cut -c1-330 gpurun_out/${T}_powerprobe_tunes.jsonl
tail -3 gpurun_out/${T}_powerprobe.err
tools/gpu_job.sh $T list cfg4
