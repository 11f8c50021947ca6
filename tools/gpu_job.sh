#!/bin/bash
# One parameterised GPU job (replaces the numbered one-shot scripts of round 1).
#   tools/gpu_job.sh TAG MODE [CONFIG] [extra bench args]
# MODE: tests    -- pytest -m gpu (whole suite) -> gpurun_out/TAG_pytest.txt
#       quick    -- pytest of the files in $FILES -> gpurun_out/TAG_pytest.txt
#       bench    -- python bench.py --config CONFIG -> gpurun_out/TAG_bench_CONFIG.json
#       list     -- ncu launch list of a short bench run -> gpurun_out/TAG_CONFIG_launches.csv
#       full     -- ncu --set full of the stream kernel -> gpurun_out/TAG_CONFIG_stream.ncu-rep
#       smoke    -- __graft_entry__.smoke()
# Run under gpurun from the repository root, e.g.
#   gpurun --timeout 900 -- 'tools/gpu_job.sh r02a bench cfg4; tools/gpu_job.sh r02a list cfg4'
set -u
TAG=$1; MODE=$2; CFG=${3:-cfg4}; shift; shift; shift || true
mkdir -p gpurun_out
KREGEX='paged_attention|reshape_and_cache|prefill'
case $MODE in
  tests) python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/${TAG}_pytest.txt; cat gpurun_out/${TAG}_pytest.txt ;;
  quick) python -m pytest ${FILES} -m gpu -x -q 2>&1 | tail -25 > gpurun_out/${TAG}_pytest_quick.txt; cat gpurun_out/${TAG}_pytest_quick.txt ;;
  bench) python bench.py --config $CFG "$@" > gpurun_out/${TAG}_bench_${CFG}.json 2> gpurun_out/${TAG}_bench_${CFG}.err; echo "bench rc=$?"; tail -3 gpurun_out/${TAG}_bench_${CFG}.err; cut -c1-1500 gpurun_out/${TAG}_bench_${CFG}.json ;;
  list)  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"$KREGEX" -c 60 --csv \
           --log-file gpurun_out/${TAG}_${CFG}_launches.csv python bench.py --config $CFG --steps 3 --warmup 3 \
           --no-cpu-baseline --no-check --sustain 0 --no-side-configs "$@" > gpurun_out/${TAG}_${CFG}_ncu_list.log 2>&1; echo "list rc=$?"
         tail -6 gpurun_out/${TAG}_${CFG}_launches.csv | awk -F'","' '{print $5, $NF}' ;;
  full)  ncu --set full --clock-control none --import-source on -k regex:'paged_attention_stream_kernel' -s 4 -c 1 -f \
           -o gpurun_out/${TAG}_${CFG}_stream python bench.py --config $CFG --steps 3 --warmup 3 \
           --no-cpu-baseline --no-check --sustain 0 --no-side-configs "$@" > gpurun_out/${TAG}_${CFG}_ncu_full.log 2>&1; echo "full rc=$?" ;;
  smoke) python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3 ;;
  *) echo "unknown mode $MODE"; exit 2 ;;
esac
