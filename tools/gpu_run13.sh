# 1 GPU: full gpu suite, small-job probe, cfg2 + cfg1 benches
O=gpurun_out; T=${TAG:-r01l}; mkdir -p $O
timeout 1800 python -m pytest tests -x -q -m gpu > $O/${T}_gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -4 $O/${T}_gpu_tests.log
python tools/small_job_probe.py 1 2>&1 | tee $O/${T}_small_probe.txt | grep -v "N      8\|N    100"
for spec in "2 100" "1 2000"; do set -- $spec
  timeout 900 python bench.py --cfg $1 --steps $2 --warmup 3 --no-cpu > $O/${T}_bench_cfg$1.json 2> $O/${T}_bench_cfg$1.err; echo "cfg$1 rc=$?"; tail -2 $O/${T}_bench_cfg$1.err
  python -c "import json;d=json.load(open('$O/${T}_bench_cfg$1.json'));print('cfg$1',d['value'],d['ms_per_step'],d['param_evals_per_s'],d['e2e']['value'],d['e2e']['ms_per_step'],d['issue']['frac'],d['other_mode']['value'],d['other_mode']['ms_per_step'])"
done
