# C++ node (reference-shaped threads: one parameter thread per stream + one gpuWorker thread) at the
# reference's own scale: param-evals/s against the number of stream slots
O=gpurun_out; mkdir -p $O /tmp/nt
python - <<'PY'
import numpy as np, sys
sys.path.insert(0, ".")
from gsmfretda_b200 import synth
w = synth.workload(1); b = w.bursts()
synth.save_bin(b, "/tmp/nt/b.bin")
rng = np.random.default_rng(1)
with open("/tmp/nt/p.txt", "w") as f:
    for i in range(60000):
        v = [rng.uniform(0.1, 0.45), rng.uniform(0.55, 0.9), rng.uniform(50, 2000), rng.uniform(50, 2000), rng.uniform(1, 4), rng.uniform(1, 4)]
        f.write("2 " + " ".join(repr(float(np.float32(x))) for x in v) + "\n")
PY
{ echo "# one launch per stream (the reference's gpuWorker loop)"
for S in 1 4 16 32; do ./gsmfretda_b200/host/local_node /tmp/nt/b.bin /tmp/nt/p.txt /tmp/nt/o.txt $S 4 7 | tail -1; done
echo "# coalesced launches (gpuWorker::set_coalesce(2, target, linger_us): vectors that are ready wait up to linger_us for company while the GPU is busy, then share one launch)"
for S in 16 32 64; do ./gsmfretda_b200/host/local_node /tmp/nt/b.bin /tmp/nt/p.txt /tmp/nt/o2.txt $S 4 7 2 | tail -2; done
echo "# fully coalesced (target 16, linger 100 us): launches drop 16x, the rate does not move: the hand-off between the threads bounds it"
./gsmfretda_b200/host/local_node /tmp/nt/b.bin /tmp/nt/p.txt /tmp/nt/o3.txt 32 4 7 2 0 16 100 | tail -2
echo "# batched requests: every parameter thread asks for P vectors at a time, ONE launch per request (set_params_batch = the batched p_ga)"
for PB in 8 32 128; do for S in 4 16; do echo "batch $PB, $S streams"; ./gsmfretda_b200/host/local_node /tmp/nt/b.bin /tmp/nt/p.txt /tmp/nt/o4.txt $S 4 7 0 0 16 60 $PB | tail -1; done; done
cmp /tmp/nt/o.txt /tmp/nt/o4.txt > /dev/null || { cut -d' ' -f1,3- /tmp/nt/o.txt | sort -n > /tmp/nt/a; cut -d' ' -f1,3- /tmp/nt/o4.txt | sort -n > /tmp/nt/c; cmp /tmp/nt/a /tmp/nt/c && echo "# batched output: identical chi-square and histogram per vector"; }
cmp /tmp/nt/o.txt /tmp/nt/o2.txt > /dev/null && echo "# outputs of the two modes: identical files" || { cut -d' ' -f1,3- /tmp/nt/o.txt | sort -n > /tmp/nt/a; cut -d' ' -f1,3- /tmp/nt/o2.txt | sort -n > /tmp/nt/b; cmp /tmp/nt/a /tmp/nt/b && echo "# outputs of the two modes: identical chi-square and histogram per vector (stream ids differ)"; }
} 2>&1 | tee $O/r02_node_throughput.txt
