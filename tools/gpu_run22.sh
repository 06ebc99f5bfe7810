O=gpurun_out; T=${TAG:-r01o}; mkdir -p $O
for spec in "5 1250000" "3 0"; do set -- $spec
extra=""; [ "$2" != "0" ] && extra="--bursts $2"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pda_eval -s 3 -c 1 -o $O/${T}_full_cfg$1 -f python bench.py --cfg $1 $extra --steps 3 --warmup 3 --no-cpu > $O/${T}_ncu_full_cfg$1.log 2>&1; echo "ncu cfg$1 rc=$?"
done
