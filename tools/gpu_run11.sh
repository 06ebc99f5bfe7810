# 1 GPU: the other configs for the record (cfg1 fixture-sized R=4, cfg3, cfg5 shard of 1.25e6 bursts)
O=gpurun_out; T=${TAG:-r01j}; mkdir -p $O
for spec in "1 2000 0" "3 30 0" "5 30 1250000"; do set -- $spec
  extra=""; [ "$3" != "0" ] && extra="--bursts $3"
  timeout 900 python bench.py --cfg $1 --steps $2 --warmup 3 --no-cpu $extra > $O/${T}_bench_cfg$1.json 2> $O/${T}_bench_cfg$1.err; echo "cfg$1 rc=$?"; tail -2 $O/${T}_bench_cfg$1.err
  python -c "import json;d=json.load(open('$O/${T}_bench_cfg$1.json'));print('cfg$1',d['value'],d['ms_per_step'],d['param_evals_per_s'],d['e2e']['value'],d['e2e']['ms_per_step'],d['issue']['frac'],d['other_mode']['value'],d['roofline']['achieved'])"
done
