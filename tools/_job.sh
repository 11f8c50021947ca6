set -u
mkdir -p gpurun_out
T=r02m
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -25 > gpurun_out/${T}_pytest.txt; tail -25 gpurun_out/${T}_pytest.txt | cut -c1-200
for t in 20 21 22 23 24 100 108; do
  B200PA_STREAM_TUNE=$t timeout 300 python tools/kbench.py --configs cfg4,cfg3 --ops fstep,flash_nhd --iters 10 --sustain 3 >> gpurun_out/${T}_kbench_block.jsonl 2>> gpurun_out/${T}_kbench.err
done
for t in 20 21 22 23 24 100 110; do
  B200PA_STREAM_TUNE=$t timeout 300 python tools/kbench.py --configs cfg2 --ops fstep,flash_nhd --iters 10 --sustain 2 >> gpurun_out/${T}_kbench_block.jsonl 2>> gpurun_out/${T}_kbench.err
done
python - <<'PY'
import json
for l in open('gpurun_out/r02m_kbench_block.jsonl'):
    d=json.loads(l); print(d['config'], d['op'], 'tune', d['tune'], d['gbs'], d.get('sustained_first_third_gbs'), d.get('sustained_last_third_gbs'))
PY
tail -3 gpurun_out/${T}_kbench.err
