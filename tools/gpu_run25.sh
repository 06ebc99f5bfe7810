O=gpurun_out; T=${TAG:-r01r}; mkdir -p $O
for n in 8 4; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700+n)) bench.py --gpus $n --steps 300 --warmup 5 > $O/${T}_bench_cfg2_n$n.json 2> $O/${T}_bench_cfg2_n$n.err; echo "n$n rc=$? lines=$(wc -l < $O/${T}_bench_cfg2_n$n.json)"
python -c "import json;d=json.load(open('$O/${T}_bench_cfg2_n$n.json'));print($n,d['value'],d['ms_per_step'],d['e2e']['value'],d['gpu_launches'],d['clocks'])"
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29790 bench.py --impl reference --gpus 8 --steps 20 --warmup 3 > $O/${T}_ref_n8.json 2> $O/${T}_ref_n8.err; echo "ref arm n8 rc=$? lines=$(wc -l < $O/${T}_ref_n8.json)"; head -c 200 $O/${T}_ref_n8.json
