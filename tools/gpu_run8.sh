O=gpurun_out; T=${TAG:-r01i}; mkdir -p $O
timeout 600 python bench.py --steps 100 > $O/${T}_bench_cfg2.json 2> $O/${T}_bench_cfg2.err; echo "cfg2 rc=$?"
python -c "import json;d=json.load(open('$O/${T}_bench_cfg2.json'));print('cfg2',d['value'],d['ms_per_step'],d['e2e']['value'],d['issue'],d['other_mode'],d['cpu_baseline']['value'])"
timeout 600 python bench.py --cfg 4 --steps 5 --warmup 3 --no-cpu > $O/${T}_bench_cfg4.json 2> $O/${T}_bench_cfg4.err; echo "cfg4 rc=$?"; tail -3 $O/${T}_bench_cfg4.err
python -c "import json;d=json.load(open('$O/${T}_bench_cfg4.json'));print('cfg4',d['value'],d['ms_per_step'],d['param_evals_per_s'],d['e2e'],d['other_mode'])"
