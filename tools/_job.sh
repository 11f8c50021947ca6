# scratch job of the current round (rewritten per call): gpurun -- 'bash tools/_job.sh'
set -u
mkdir -p gpurun_out
T=r03e
python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/${T}_pytest_all.txt
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2 | tee gpurun_out/${T}_smoke.txt
python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo bench rc=$?; tail -3 gpurun_out/${T}_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err; echo ref rc=$?; cut -c1-300 gpurun_out/${T}_bench_reference.json
python - <<PY
import json
l=json.loads(open("gpurun_out/${T}_bench.json").read().strip().splitlines()[-1])
print(l["value"], l["ms_per_step"], l["e2e"]["value"], l["parity_status"], l["ref_cuda"].get("speedup"), l["roofline"]["frac"], l["clocks"])
for k,v in l.get("configs",{}).items(): print(k, v.get("gbs"), v.get("ms_per_step"), v.get("parity_status"), v.get("ref_cuda",{}).get("speedup"), v.get("error"))
print(l.get("sustained",{}).get("first_vs_last_third"))
PY
