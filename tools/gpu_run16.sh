# 8 GPUs: config 5 as BASELINE.json states it — 1e7 bursts (4 states, background) split across 8 B200
O=gpurun_out; T=${TAG:-r01m}; mkdir -p $O
for coll in peer nccl; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29655 bench.py --gpus 8 --cfg 5 --bursts 1250000 --steps 50 --warmup 5 --no-cpu --collective $coll > $O/${T}_bench_cfg5_n8_$coll.json 2> $O/${T}_bench_cfg5_n8_$coll.err; echo "cfg5 n8 $coll rc=$?"; tail -2 $O/${T}_bench_cfg5_n8_$coll.err
python - $O/${T}_bench_cfg5_n8_$coll.json <<'PY'
import json,sys
for ln in open(sys.argv[1]):
    if ln.startswith('{'):
        d=json.loads(ln); print('  ', d['n_gpus'], d['value'], d['ms_per_step'], d['param_evals_per_s'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'], 'normal', d['other_mode']['value'])
PY
done
