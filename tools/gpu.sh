#!/bin/bash
# tools/gpu.sh — the one parameterised GPU-box script (run through gpurun).
#   usage: TAG=r02a bash tools/gpu.sh <job> [job ...]
#   jobs: tests | bench:<cfg>[:<mode>[:<extra bench args>]] | launches | ncu:<cfg>[:<mode>[:<bursts>]] | variants:<cfg>:<lib,lib,...>
O=gpurun_out; T=${TAG:-x}; mkdir -p $O
for job in "$@"; do
  IFS=: read -r kind a1 a2 a3 <<< "$job"
  case $kind in
    tests)
      timeout 1500 python -m pytest tests -m gpu -x -q > $O/${T}_gpu_tests.log 2>&1; echo "tests rc=$? $(tail -1 $O/${T}_gpu_tests.log)";;
    bench)
      M=${a2:-replay}
      timeout 600 python bench.py --cfg $a1 --only main --mode $M --no-cpu $a3 > $O/${T}_bench_cfg${a1}_$M.json 2> $O/${T}_bench_cfg${a1}_$M.err; echo "bench cfg$a1 $M rc=$?"
      python - <<PY
import json
try:
    d=json.loads(open('$O/${T}_bench_cfg${a1}_$M.json').read().strip().splitlines()[-1])
    print('cfg$a1 $M value %.4g ms %.4f e2e %.4g' % (d['value'], d['ms_per_step'], d['e2e']['value']))
except Exception as e: print('parse failed', e)
PY
      ;;
    launches)
      timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches.csv python bench.py --steps 20 --warmup 3 --no-cpu --sustained 0 > $O/${T}_ncu_list.log 2>&1; echo "ncu list rc=$?";;
    ncu)
      M=${a2:-replay}; B=""; [ -n "$a3" ] && B="--bursts $a3"
      timeout 900 ncu --set full --clock-control none --import-source on -k regex:pda_eval -s 3 -c 1 -o $O/${T}_full_cfg${a1}_$M -f python bench.py --cfg $a1 --only main $B --steps 3 --warmup 3 --no-cpu --sustained 0 --mode $M > $O/${T}_ncu_full_cfg${a1}_$M.log 2>&1; echo "ncu cfg$a1 $M rc=$?";;
    variants)
      bash tools/variant_bench.sh --cfg $a1 --steps 100 --warmup 5 $a3 -- $(echo $a2 | tr ',' ' ') | tee -a $O/${T}_variants_cfg$a1.txt;;
    *) echo "unknown job $job";;
  esac
done
