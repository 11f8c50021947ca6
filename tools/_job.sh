set -u
mkdir -p gpurun_out
T=r02p
nvidia-smi -L | head -8
timeout 600 python -m pytest tests/test_gpu_multi_rank.py -m gpu -q 2>&1 | tail -5 | tee gpurun_out/${T}_pytest_multi_rank.txt
P=29500
for N in 8 4 2; do
  P=$((P+1))
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N > gpurun_out/${T}_scale_n$N.json 2> gpurun_out/${T}_scale_n$N.err
  echo "N=$N rc=$?"; tail -2 gpurun_out/${T}_scale_n$N.err | cut -c1-300
done
timeout 900 python bench.py --gpus 1 > gpurun_out/${T}_scale_n1.json 2> gpurun_out/${T}_scale_n1.err; echo "N=1 rc=$?"
python - <<'PY'
import json
base=None
for n in (1,2,4,8):
    try:
        d=json.load(open(f'gpurun_out/r02p_scale_n{n}.json'))
    except Exception as e:
        print(n,'unreadable',e); continue
    if n==1: base=d['value']
    print('N',n,'value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],'sustained',d.get('sustained',{}).get('gbs'),'parity',d['parity'].get('ok'), d['parity'].get('gathered_vs_unsharded',{}).get('ok'), 'x', round(d['value']/base,2) if base else None)
    print('   ', {k:(v['gbs'],v['ms_per_step'],v['parity'].get('ok')) for k,v in d.get('configs',{}).items()})
PY
