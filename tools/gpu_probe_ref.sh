set -x
O=gpurun_out
for spec in "1 2000 64" "1 20000 4" "1 20000 16" "2 2000 4" "2 20000 4" "1 20000 64"; do
  set -- $spec
  timeout 90 python -X faulthandler oracle/refgpu_run.py --cfg $1 --bursts $2 --repeats $3 --evals 2 --out $O/refgpu_p_$1_$2_$3.npz > $O/refgpu_p_$1_$2_$3.log 2>&1; echo "refgpu $spec rc=$?"; tail -6 $O/refgpu_p_$1_$2_$3.log
done
