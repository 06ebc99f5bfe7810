#!/bin/bash
O=gpurun_out; T=${TAG:-r02p}; mkdir -p $O
bash tools/node_throughput.sh > $O/${T}_node_throughput.log 2>&1; cp $O/r02_node_throughput.txt $O/${T}_node_throughput.txt; cat $O/${T}_node_throughput.txt
timeout 600 python -m pytest tests/test_host_cpp.py -m gpu -x -q 2>&1 | tail -2
