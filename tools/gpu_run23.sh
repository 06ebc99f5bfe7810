O=gpurun_out; T=${TAG:-r01p}; mkdir -p $O
timeout 1800 python -m pytest tests -q -m gpu -x > $O/${T}_gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -4 $O/${T}_gpu_tests.log
timeout 600 python bench.py --cfg 5 --bursts 1250000 --steps 30 --warmup 3 --no-cpu > $O/${T}_bench_cfg5.json 2> $O/${T}_bench_cfg5.err; echo "bench rc=$?"
python -c "import json;d=json.load(open('$O/${T}_bench_cfg5.json'));print('cfg5',d['value'],d['ms_per_step'],d['e2e']['value'],d['other_mode']['value'])"
timeout 600 python bench.py --steps 100 --no-cpu > $O/${T}_bench_cfg2.json 2> $O/${T}_bench_cfg2.err; echo "bench rc=$?"
python -c "import json;d=json.load(open('$O/${T}_bench_cfg2.json'));print('cfg2',d['value'],d['ms_per_step'],d['e2e']['value'],d['other_mode']['value'])"
