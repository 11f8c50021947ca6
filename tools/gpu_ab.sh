#!/bin/bash
# Same-box A/B of two builds of the package: _ab/old (a copy of the package built from an
# earlier commit) vs the tree.  OPS / SPECS / REPS select what is timed.
# Alternating old,new x REPS per shard, 20 launches each (median + best per process), no profiler.
REPS=${REPS:-3}
OPS=${OPS:-v2}
SPECS=${SPECS:-"cfg1 1;cfg5 8;cfg4 8;cfg2 1;cfg3 1;cfg5 1;cfg4 1"}
mkdir -p gpurun_out; rm -f gpurun_out/ab.jsonl gpurun_out/ab.err
IFS=";" read -ra SPEC_LIST <<< "$SPECS"
for spec in "${SPEC_LIST[@]}"; do
  IFS=" " read -r c t <<< "$spec"; set -- $c $t
  for rep in $(seq $REPS); do
    B200PA_AB_LABEL=old B200PA_PKG_ROOT=$PWD/_ab/old python tools/kbench.py --configs $1 --tp $2 --ops $OPS --iters 20 >> gpurun_out/ab.jsonl 2>> gpurun_out/ab.err
    B200PA_AB_LABEL=new python tools/kbench.py --configs $1 --tp $2 --ops $OPS --iters 20 >> gpurun_out/ab.jsonl 2>> gpurun_out/ab.err
  done
done
python - <<'PY'
import json, collections
rows = [json.loads(l) for l in open("gpurun_out/ab.jsonl")]
d = collections.defaultdict(list)
for r in rows:
    d[(r["config"] + " " + r["op"], r["tp"], r["build"])].append(r["ms_median"] * 1e3)
for (cfg, tp) in sorted({(k[0], k[1]) for k in d}):
    o, n = d[(cfg, tp, "old")], d[(cfg, tp, "new")]
    fmt = lambda v: "[" + " ".join(f"{x:.1f}" for x in v) + "]"
    mo, mn = sorted(o)[len(o)//2], sorted(n)[len(n)//2]
    print(f"{cfg} tp{tp}: old us {fmt(o)} new us {fmt(n)}  median old {mo:.1f} new {mn:.1f}  delta {100*(mn-mo)/mo:+.1f}%")
PY
tail -3 gpurun_out/ab.err
