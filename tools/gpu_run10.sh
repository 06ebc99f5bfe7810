# 2 GPUs: peer-exchange tests (IPC across devices), then burst-split bench peer vs nccl, population bench
O=gpurun_out; T=${TAG:-r01j}; mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv,noheader; nvidia-smi topo -m | head -6
timeout 600 python -m pytest tests/test_gpu_peer.py -x -q -m gpu > $O/${T}_peer_tests_2gpu.log 2>&1; echo "peer tests rc=$?"; tail -5 $O/${T}_peer_tests_2gpu.log
for coll in peer nccl; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 300 --warmup 5 --no-cpu --collective $coll > $O/${T}_bench_cfg2_n2_$coll.json 2> $O/${T}_bench_cfg2_n2_$coll.err; echo "bench n2 $coll rc=$?"; tail -3 $O/${T}_bench_cfg2_n2_$coll.err
python -c "import json;d=json.load(open('$O/${T}_bench_cfg2_n2_$coll.json'));print('$coll',d['value'],d['ms_per_step'],d['e2e']['value'],d['gpu_launches'],d['config']['sharding'])"
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --cfg 4 --steps 5 --warmup 3 --no-cpu > $O/${T}_bench_cfg4_n2.json 2> $O/${T}_bench_cfg4_n2.err; echo "cfg4 n2 rc=$?"
python -c "import json;d=json.load(open('$O/${T}_bench_cfg4_n2.json'));print('cfg4 n2',d['value'],d['ms_per_step'],d['param_evals_per_s'])"
