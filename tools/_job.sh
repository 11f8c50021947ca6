set -u
mkdir -p gpurun_out
T=r02j
python -m pytest tests -m gpu -q 2>&1 | tail -12 > gpurun_out/${T}_pytest.txt; tail -6 gpurun_out/${T}_pytest.txt
python tools/opbench.py 2> gpurun_out/${T}_opbench.err | grep reshape_and_cache > gpurun_out/${T}_opbench_cache.jsonl; grep '"reshape_and_cache ' gpurun_out/${T}_opbench_cache.jsonl | grep 16384 | cut -c1-200
for t in 0 1 2; do
  python tools/shapebench.py --only d192_bf16,d96_fp8,d64_bf16_gqa8,d80_fp8 --tune $t --sustain 2 --stream-only >> gpurun_out/${T}_shapebench_tunes.jsonl 2>> gpurun_out/${T}_shapebench.err
done
python tools/shapebench.py --sustain 2 --stream-only > gpurun_out/${T}_shapebench_all.jsonl 2>> gpurun_out/${T}_shapebench.err
python - <<'PY'
import json
for f in ('gpurun_out/r02j_shapebench_tunes.jsonl','gpurun_out/r02j_shapebench_all.jsonl'):
    print(f)
    for l in open(f):
        d=json.loads(l)
        if 'shape' in d: print(d['shape'], d['tune'], d['gbs'], d.get('sustained_first_third_gbs'), d.get('sustained_last_third_gbs'))
PY
for t in 0 8; do
  for c in cfg4 cfg5; do
    timeout 120 python tools/powerprobe.py --config $c --tp 8 --seconds 3 --tune $t >> gpurun_out/${T}_powerprobe_tp8.jsonl 2>> gpurun_out/${T}_powerprobe.err
    sleep 2
  done
done
python - <<'PY'
import json
for l in open('gpurun_out/r02j_powerprobe_tp8.jsonl'):
    d=json.loads(l); print(d['config'], 'tp', d['tp'], 'tune', d['tune'], d['burst_gbs_10_steps'], d['first_300ms_gbs'], d['last_third_gbs'], d['sm_mhz_last_third'], d['power_w_last_third'])
PY
python bench.py > gpurun_out/${T}_bench_cfg4.json 2> gpurun_out/${T}_bench_cfg4.err; echo "bench rc=$?"; tail -2 gpurun_out/${T}_bench_cfg4.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02j_bench_cfg4.json'))
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],'sustained',d['sustained']['gbs'],'frac',d['roofline']['frac'],'parity',d['parity']['ok'])
print({k:(v['gbs'],v['ms_per_step'],v['parity'].get('ok')) for k,v in d['configs'].items()})
PY
