#!/bin/bash
# scratch job for one gpurun call (rewritten per call; tools/gpu.sh holds the reusable jobs)
O=gpurun_out; T=${TAG:-r02j}; mkdir -p $O
./tools/log2_probe.bin > $O/${T}_log2_probe.txt 2>&1; cat $O/${T}_log2_probe.txt | grep -c "off by one"
timeout 1700 python -m pytest tests -m gpu -x -q > $O/${T}_gpu_tests.log 2>&1; echo "tests rc=$? $(tail -1 $O/${T}_gpu_tests.log)"
timeout 900 python bench.py > $O/${T}_bench_default.json 2> $O/${T}_bench_default.err; echo "bench rc=$?"; tail -8 $O/${T}_bench_default.err
python - <<PY
import json
d=json.loads(open('$O/${T}_bench_default.json').read().strip().splitlines()[-1])
print('main %.4g ms %.4f e2e %.4g frac %.3f' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac']))
for k,v in d.get('configs',{}).items(): print(k, '%.4g ms %.4f e2e %.4g pe/s %.4g frac %.3f' % (v['value'], v['ms_per_step'], v['e2e']['value'], v['param_evals_per_s'], v['roofline']['frac']))
print(json.dumps(d.get('cpu_baseline'))[:3000])
PY
