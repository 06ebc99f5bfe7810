O=gpurun_out; T=${TAG:-r01o}; mkdir -p $O
timeout 1800 python -m pytest tests -q -m gpu > $O/${T}_gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -5 $O/${T}_gpu_tests.log
for spec in "2 100" "1 2000"; do set -- $spec
timeout 600 python bench.py --cfg $1 --mode reference --steps $2 --no-cpu > $O/${T}_bench_cfg$1_reference.json 2> $O/${T}_bench_cfg$1_reference.err; echo "bench rc=$?"; tail -2 $O/${T}_bench_cfg$1_reference.err
python -c "import json;d=json.load(open('$O/${T}_bench_cfg$1_reference.json'));print('cfg$1 reference mode',d['value'],d['ms_per_step'],d['e2e']['value'],d['other_mode'])"
done
