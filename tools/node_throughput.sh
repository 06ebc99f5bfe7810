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
    for i in range(20000):
        v = [rng.uniform(0.1, 0.45), rng.uniform(0.55, 0.9), rng.uniform(50, 2000), rng.uniform(50, 2000), rng.uniform(1, 4), rng.uniform(1, 4)]
        f.write("2 " + " ".join(repr(float(np.float32(x))) for x in v) + "\n")
PY
for S in 1 2 4 8 16 32; do ./gsmfretda_b200/host/local_node /tmp/nt/b.bin /tmp/nt/p.txt /tmp/nt/o.txt $S 4 7 | tail -1; done 2>&1 | tee $O/r01q_node_throughput.txt
