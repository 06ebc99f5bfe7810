# 1 GPU: ncu evidence for the current kernel — launch list of the bench command, full captures of both modes
O=gpurun_out; T=${TAG:-r01l}; mkdir -p $O
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu > $O/${T}_bench_short.json 2> $O/${T}_bench_short.err; echo "bench rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches.csv python bench.py --steps 20 --warmup 3 --no-cpu > $O/${T}_ncu_list.log 2>&1; echo "ncu list rc=$?"
for M in replay normal; do
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pda_eval -s 3 -c 1 -o $O/${T}_full_$M -f python bench.py --steps 3 --warmup 3 --no-cpu --mode $M > $O/${T}_ncu_full_$M.log 2>&1; echo "ncu $M rc=$?"
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pda_eval -s 3 -c 1 -o $O/${T}_full_cfg1 -f python bench.py --cfg 1 --steps 3 --warmup 3 --no-cpu > $O/${T}_ncu_full_cfg1.log 2>&1; echo "ncu cfg1 rc=$?"
ls -la $O/*.ncu-rep
