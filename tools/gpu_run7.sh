V="default $(ls variants/*.so 2>/dev/null | tr '\n' ' ')"
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
echo "== replay cfg2"; bash tools/variant_bench.sh --steps 50 -- $V
echo "== normal cfg2"; bash tools/variant_bench.sh --steps 50 --mode normal -- $V
echo "== replay cfg5"; bash tools/variant_bench.sh --steps 5 --warmup 3 --cfg 5 -- $V
