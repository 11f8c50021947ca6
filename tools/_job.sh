# scratch job of the current round (rewritten per call): gpurun -- 'bash tools/_job.sh'
set -u
mkdir -p gpurun_out
T=r03c
python -m pytest tests/test_gpu_prefill.py tests/test_gpu_backend_multi_query.py -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/${T}_pytest_prefill.txt
python -m pytest tests/test_gpu_attention.py -m gpu -x -q -k "scratch or opcheck or schema or forward_decode or backend" 2>&1 | tail -6 | tee gpurun_out/${T}_pytest_v2.txt
python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo bench rc=$?; tail -3 gpurun_out/${T}_bench.err
python - <<PY
import json
l=json.loads(open("gpurun_out/${T}_bench.json").read().strip().splitlines()[-1])
print(l["value"], l["ms_per_step"], l["e2e"]["value"], l["parity_status"], l["ref_cuda"], l["roofline"]["frac"])
for k,v in l.get("configs",{}).items(): print(k, v.get("gbs"), v.get("ms_per_step"), v.get("parity_status"), v.get("ref_cuda"), v.get("error"))
print(l.get("sustained",{}).get("first_vs_last_third"))
PY
