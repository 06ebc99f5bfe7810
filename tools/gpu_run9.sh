# 1 GPU: peer-exchange tests first, then the whole GPU suite, then the default bench
O=gpurun_out; T=${TAG:-r01j}; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
timeout 900 python -m pytest tests/test_gpu_peer.py -x -q -m gpu > $O/${T}_peer_tests.log 2>&1; echo "peer tests rc=$?"; tail -15 $O/${T}_peer_tests.log
timeout 1800 python -m pytest tests -q -m gpu --deselect tests/test_gpu_peer.py > $O/${T}_gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -5 $O/${T}_gpu_tests.log
timeout 600 python bench.py > $O/${T}_bench_cfg2.json 2> $O/${T}_bench_cfg2.err; echo "bench rc=$?"; tail -3 $O/${T}_bench_cfg2.err
python -c "import json;d=json.load(open('$O/${T}_bench_cfg2.json'));print('cfg2',d['value'],d['ms_per_step'],d['e2e']['value'],d['issue']['frac'],d['other_mode']['value'],d['cpu_baseline']['value'])"
