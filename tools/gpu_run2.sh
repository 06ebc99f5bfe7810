set -x
O=gpurun_out
nvidia-smi -L; nproc
timeout 900 python -m pytest tests -m gpu -x -q > $O/r01b_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 $O/r01b_pytest_gpu.log
timeout 300 python bench.py --no-cpu > $O/r01b_bench_replay.json 2> $O/r01b_bench_replay.err; echo "bench rc=$?"; cat $O/r01b_bench_replay.json
timeout 300 python bench.py --no-cpu --mode normal > $O/r01b_bench_normal.json 2> $O/r01b_bench_normal.err; echo "bench rc=$?"; cat $O/r01b_bench_normal.json
for c in 1 3 5; do timeout 300 python bench.py --cfg $c --steps 20 --warmup 3 --no-cpu --mode normal > $O/r01b_bench_cfg${c}_normal.json 2>$O/r01b_bench_cfg${c}_normal.err; cat $O/r01b_bench_cfg${c}_normal.json; done
timeout 400 ncu --set full --clock-control none --import-source on -k regex:pda_eval -s 3 -c 1 -o $O/r01b_full_normal -f python bench.py --steps 2 --warmup 3 --no-cpu --mode normal > $O/r01b_ncu_full_normal.log 2>&1; echo "ncu rc=$?"
# the reference's own class mc on this GPU: tiny first, each bounded
timeout 120 python -X faulthandler oracle/refgpu_run.py --cfg 1 --bursts 2000 --repeats 4 --evals 1 --out $O/refgpu_tiny.npz > $O/refgpu_tiny.log 2>&1; echo "refgpu tiny rc=$?"; tail -5 $O/refgpu_tiny.log
timeout 200 python -X faulthandler oracle/refgpu_run.py --cfg 2 --bursts 20000 --repeats 64 --evals 2 --time 5 --out $O/refgpu_cfg2.npz > $O/refgpu_cfg2.log 2>&1; echo "refgpu cfg2 rc=$?"; tail -5 $O/refgpu_cfg2.log
