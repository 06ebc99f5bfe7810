O=gpurun_out; T=${TAG:-r01o}; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_reference_mode.py tests/test_gpu_reference_gpu.py -x -q -m gpu > $O/${T}_refmode_tests.log 2>&1; echo "ref-mode tests rc=$?"; tail -15 $O/${T}_refmode_tests.log
