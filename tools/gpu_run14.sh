# 8 GPUs: scaling of both sharding modes (what the driver's SCALE run does), peer exchange at 4 and 8 ranks
O=gpurun_out; T=${TAG:-r01l}; mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
run() { # n cfg steps extra
  n=$1; cfg=$2; steps=$3; shift 3
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n + cfg)) bench.py --gpus $n --cfg $cfg --steps $steps --warmup 5 --no-cpu "$@" > $O/${T}_bench_cfg${cfg}_n${n}$TAG2.json 2> $O/${T}_bench_cfg${cfg}_n${n}$TAG2.err; echo "cfg$cfg n$n $* rc=$?"
  python - $O/${T}_bench_cfg${cfg}_n${n}$TAG2.json <<'PY'
import json,sys
for ln in open(sys.argv[1]):
    if ln.startswith('{'):
        d=json.loads(ln); print('  ', d['n_gpus'], d['value'], d['ms_per_step'], d['param_evals_per_s'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'], d['config']['sharding'][:40])
PY
}
TAG2=""; run 8 2 300; run 4 2 300
TAG2="_nccl"; run 8 2 300 --collective nccl
TAG2=""; run 8 4 5; run 4 4 5
timeout 300 python -m pytest tests/test_gpu_peer.py -x -q -m gpu 2>&1 | tail -2
