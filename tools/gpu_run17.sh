# 1 GPU: this library against the reference's own class mc running on the same B200 (oracle/_ref/libref_gpu.so)
O=gpurun_out; T=${TAG:-r01m}; mkdir -p $O
for m in 0 1; do
timeout 600 python tools/refgpu_compare.py --cfgs 1,2,3,5 --bursts 60000 --repeats 16 --mode $m --timeout 150 > $O/${T}_refgpu_cmp_mode$m.jsonl 2> $O/${T}_refgpu_cmp_mode$m.err; echo "cmp mode $m rc=$?"; tail -3 $O/${T}_refgpu_cmp_mode$m.err
python - $O/${T}_refgpu_cmp_mode$m.jsonl <<'PY'
import json,sys
for ln in open(sys.argv[1]):
    d=json.loads(ln); print({k:d[k] for k in ('cfg','R','mode','ks_ours_vs_ref','ks_ref_vs_ref','ks_ours_vs_ours','chi2_ours','chi2_ref','chi2_rel_diff','nan_ref')})
PY
done
for spec in "2 100000 4" "2 100000 64" "1 7733 4"; do set -- $spec
timeout 200 python oracle/refgpu_run.py --cfg $1 --bursts $2 --repeats $3 --evals 1 --time 10 > $O/${T}_refgpu_time_cfg$1_R$3.json 2> $O/${T}_refgpu_time_cfg$1_R$3.err; echo "time cfg$1 R$3 rc=$?"; cat $O/${T}_refgpu_time_cfg$1_R$3.json
done
