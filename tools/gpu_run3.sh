set -x
O=gpurun_out
timeout 600 python tools/normal_diag.py > $O/normal_diag.jsonl 2> $O/normal_diag.err; echo "diag rc=$?"; cat $O/normal_diag.jsonl; tail -3 $O/normal_diag.err
for m in 0 1; do
timeout 500 python tools/refgpu_compare.py --cfgs 1,2,3,5 --bursts 60000 --repeats 16 --mode $m --timeout 100 > $O/refgpu_cmp_mode$m.jsonl 2> $O/refgpu_cmp_mode$m.err; echo "cmp rc=$?"; cat $O/refgpu_cmp_mode$m.jsonl; tail -3 $O/refgpu_cmp_mode$m.err
done
timeout 100 python oracle/refgpu_run.py --cfg 2 --bursts 100000 --repeats 16 --evals 1 --time 10 > $O/refgpu_time_cfg2_R16.json 2> $O/refgpu_time_cfg2_R16.err; cat $O/refgpu_time_cfg2_R16.json
timeout 100 python oracle/refgpu_run.py --cfg 2 --bursts 100000 --repeats 4 --evals 1 --time 10 > $O/refgpu_time_cfg2_R4.json 2> $O/refgpu_time_cfg2_R4.err; cat $O/refgpu_time_cfg2_R4.json
for R in 32 128; do timeout 60 python oracle/refgpu_run.py --cfg 1 --bursts 2000 --repeats $R --evals 1 > $O/refgpu_R$R.log 2>&1; echo "refgpu R=$R rc=$?"; done
