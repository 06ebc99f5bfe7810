#!/bin/bash
# usage: tools/variant_bench.sh [bench args] -- lib1.so lib2.so ...   ("default" = in-tree library)
args=()
while [ $# -gt 0 ] && [ "$1" != "--" ]; do args+=("$1"); shift; done
shift
for v in "$@"; do
  if [ "$v" = "default" ]; then unset GSMFRET_LIB; else export GSMFRET_LIB="$PWD/$v"; fi
  out=$(python bench.py --no-cpu --sustained 0 "${args[@]}" 2>/tmp/vb.err | tail -1)
  echo "$out" | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read())
    print('$v', 'value %.3f G/s' % (d['value']/1e9), 'ms %.4f' % d['ms_per_step'], 'e2e %.3f G/s' % (d['e2e']['value']/1e9))
except Exception as e:
    print('$v', 'FAILED', e); print(open('/tmp/vb.err').read()[-800:])
"
done
