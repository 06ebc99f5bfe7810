O=gpurun_out; T=${TAG:-r01r}; mkdir -p $O
bash tools/node_throughput.sh | tail -6
timeout 1800 python -m pytest tests -q -m gpu -x > $O/${T}_gpu_tests.log 2>&1; echo "gpu tests rc=$?"; tail -3 $O/${T}_gpu_tests.log
