# usage: TAG=x MODE=replay|normal bash tools/gpu_ncu_one.sh   -- one full ncu capture of the evaluation kernel
O=gpurun_out; T=${TAG:-x}; M=${MODE:-replay}
mkdir -p $O
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu --mode $M > $O/${T}_bench_$M.json 2> $O/${T}_bench_$M.err; echo "bench rc=$?"
python -c "import json;d=json.load(open('$O/${T}_bench_$M.json'));print('$M',d['value'],d['ms_per_step'])"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pda_eval -s 3 -c 1 -o $O/${T}_full_$M -f python bench.py --steps 3 --warmup 3 --no-cpu --sustained 0 --mode $M > $O/${T}_ncu_full_$M.log 2>&1; echo "ncu rc=$?"
