set -x
nvidia-smi -L; nproc
python -m pytest tests -m gpu -x -q > gpurun_out/r01_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r01_pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/r01_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r01_smoke.log
python bench.py > gpurun_out/r01_bench.json 2> gpurun_out/r01_bench.err; echo "bench rc=$?"; cat gpurun_out/r01_bench.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r01_bench_ref.json 2>&1; cat gpurun_out/r01_bench_ref.json
for c in 1 3 5; do python bench.py --cfg $c --steps 20 --warmup 3 --no-cpu > gpurun_out/r01_bench_cfg$c.json 2>gpurun_out/r01_bench_cfg$c.err; cat gpurun_out/r01_bench_cfg$c.json; done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pda_eval -s 3 -c 1 -o gpurun_out/r01_full -f python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r01_ncu_full.log 2>&1; echo "ncu rc=$?"
