#!/bin/bash
# tools/sweep.sh — one bench line per spec, on the GPU box.  spec = lib|occ|cfg|extra bench args
#   lib: "default" or variants/libgsmfret_<name>.so ; occ: 0, 1 or - (GSM_OCC: register budget of the evaluation kernel)
#   usage: TAG=r02s bash tools/sweep.sh 'default|0|2|' 'variants/libgsmfret_x.so|1|5|--bursts 1250000 --steps 50'
O=gpurun_out; T=${TAG:-x}; mkdir -p $O
for spec in "$@"; do
  IFS='|' read -r lib occ cfg extra <<< "$spec"
  if [ "$lib" = "default" ]; then unset GSMFRET_LIB; else export GSMFRET_LIB="$PWD/$lib"; fi
  if [ "$occ" = "-" ]; then unset GSM_OCC; else export GSM_OCC=$occ; fi
  out=$(timeout 300 python bench.py --no-cpu --sustained 0 --only main --cfg $cfg --steps 100 --warmup 5 $extra 2>/tmp/sw.err | tail -1)
  echo "$out" | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read())
    print('$lib occ=$occ cfg$cfg $extra:', 'value %.3f G/s' % (d['value']/1e9), 'ms %.4f' % d['ms_per_step'], 'e2e %.3f G/s' % (d['e2e']['value']/1e9))
except Exception as e:
    print('$lib occ=$occ cfg$cfg', 'FAILED', e); print(open('/tmp/sw.err').read()[-800:])
" | tee -a $O/${T}_sweep.txt
done
