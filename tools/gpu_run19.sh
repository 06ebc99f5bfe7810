O=gpurun_out; T=${TAG:-r01n}; mkdir -p $O
timeout 1800 python -m pytest tests -q -m gpu > $O/${T}_gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -6 $O/${T}_gpu_tests.log
timeout 600 python bench.py --steps 100 --no-cpu > $O/${T}_bench_cfg2.json 2> $O/${T}_bench_cfg2.err; echo "bench rc=$?"
python -c "import json;d=json.load(open('$O/${T}_bench_cfg2.json'));print('cfg2',d['value'],d['ms_per_step'],d['e2e']['value'],d['other_mode']['value'],d['roofline'])"
