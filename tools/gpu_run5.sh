set -x
O=gpurun_out
T=${TAG:-r01e}
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
nproc
timeout 1200 python -m pytest tests -m gpu -x -q > $O/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 $O/${T}_pytest_gpu.log
timeout 600 python bench.py --steps 100 > $O/${T}_bench_replay.json 2> $O/${T}_bench_replay.err; echo "bench rc=$?"; python -c "import json;d=json.load(open('$O/${T}_bench_replay.json'));print('REPLAY',d['value'],d['ms_per_step'],d['e2e'],d['issue'],d['cpu_baseline'],d['clocks'])"
timeout 300 python bench.py --no-cpu --steps 100 --mode normal > $O/${T}_bench_normal.json 2> $O/${T}_bench_normal.err; echo "bench rc=$?"; python -c "import json;d=json.load(open('$O/${T}_bench_normal.json'));print('NORMAL',d['value'],d['ms_per_step'])"
for c in 1 3 4 5; do for m in replay normal; do timeout 300 python bench.py --cfg $c --steps 20 --warmup 3 --no-cpu --mode $m > $O/${T}_bench_cfg${c}_$m.json 2>$O/${T}_bench_cfg${c}_$m.err; python -c "import json;d=json.load(open('$O/${T}_bench_cfg${c}_$m.json'));print('cfg$c $m',d['value'],d['ms_per_step'])"; done; done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches.csv python bench.py --steps 5 --warmup 3 --no-cpu > $O/${T}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pda_eval -s 3 -c 1 -o $O/${T}_full -f python bench.py --steps 3 --warmup 3 --no-cpu > $O/${T}_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pda_eval -s 3 -c 1 -o $O/${T}_full_normal -f python bench.py --steps 3 --warmup 3 --no-cpu --mode normal > $O/${T}_ncu_full_normal.log 2>&1; echo "ncu full normal rc=$?"
ls -la $O
