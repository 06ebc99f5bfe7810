#!/bin/bash
# scratch job for one gpurun call (rewritten per call; tools/gpu.sh holds the reusable jobs)
O=gpurun_out; T=${TAG:-r02k}; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/${T}_gpu_tests.log 2>&1; echo "tests rc=$? $(tail -1 $O/${T}_gpu_tests.log)"
V="variants/libgsmfret_base.so default variants/libgsmfret_t8.so variants/libgsmfret_t16.so variants/libgsmfret_t24.so variants/libgsmfret_d12m3.so variants/libgsmfret_d6.so"
for c in 2 3; do echo "== cfg $c"; bash tools/variant_bench.sh --cfg $c --only main --steps 100 --warmup 5 -- $V; done 2>&1 | tee $O/${T}_variants.txt
echo "== cfg 5 (1.25e6)"; bash tools/variant_bench.sh --cfg 5 --bursts 1250000 --steps 50 --warmup 5 -- $V 2>&1 | tee -a $O/${T}_variants.txt
