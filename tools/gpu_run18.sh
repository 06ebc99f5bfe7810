O=gpurun_out; T=${TAG:-r01n}; mkdir -p $O
timeout 1800 python -m pytest tests -x -q -m gpu > $O/${T}_gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -4 $O/${T}_gpu_tests.log
timeout 600 python tools/refgpu_sobol_check.py 2>&1 | tee $O/${T}_refgpu_sobol_check.jsonl
