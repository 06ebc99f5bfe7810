set -x
O=gpurun_out
T=${TAG:-r01d}
timeout 900 python -m pytest tests -m gpu -x -q > $O/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 $O/${T}_pytest_gpu.log
timeout 300 python bench.py --no-cpu --steps 100 > $O/${T}_bench_replay.json 2> $O/${T}_bench_replay.err; echo "bench rc=$?"; python -c "import json;d=json.load(open('$O/${T}_bench_replay.json'));print('REPLAY',d['value'],d['ms_per_step'])"
timeout 300 python bench.py --no-cpu --steps 100 --mode normal > $O/${T}_bench_normal.json 2> $O/${T}_bench_normal.err; echo "bench rc=$?"; python -c "import json;d=json.load(open('$O/${T}_bench_normal.json'));print('NORMAL',d['value'],d['ms_per_step'])"
for c in 3 5; do for m in replay normal; do timeout 300 python bench.py --cfg $c --steps 20 --warmup 3 --no-cpu --mode $m > $O/${T}_bench_cfg${c}_$m.json 2>$O/${T}_bench_cfg${c}_$m.err; python -c "import json;d=json.load(open('$O/${T}_bench_cfg${c}_$m.json'));print('cfg$c $m',d['value'],d['ms_per_step'])"; done; done
