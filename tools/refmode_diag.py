import os, subprocess, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_py
from gsmfretda_b200 import capi, synth
cfg, nb, R = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
w = synth.workload(cfg, n_bursts=nb, R=R); b = w.bursts()
with tempfile.TemporaryDirectory() as td:
    out = os.path.join(td, "ref.npz")
    subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "refgpu_run.py"), "--cfg", str(cfg), "--bursts", str(nb), "--repeats", str(R),
                    "--evals", "2", "--out", out], check=True, timeout=200, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    gpu = np.load(out)["mcE"]
host = oracle_py.Ref().mc_kernel_sobol(b, w.n_states, w.args[0], R, evals=2)
ours = []
with capi.Mc(0, 1, keep_mce=True) as mc:
    mc.set_reSampleTimes(R); mc.set_mode(capi.GSM_MODE_REFERENCE); mc.init_bursts(b)
    for _ in range(2):
        mc.evaluate(w.n_states, w.args[0]); ours.append(mc.hmcE(0, nb * R).copy())
for ev in range(2):
    for nm, x, y in (("ours vs refgpu", ours[ev], gpu[ev]), ("ours vs host", ours[ev], host[ev]), ("host vs refgpu", host[ev], gpu[ev])):
        bad = np.nonzero(x != y)[0]
        print(ev, nm, "mismatches", bad.size, [(int(i), float(x[i]), float(y[i])) for i in bad[:5]])
