#!/usr/bin/env python
"""Structure of the per-sample mismatches between the reference GPU run and the host build of
the same sources with restated Sobol64 draws (see tools/refgpu_sobol_check.py)."""
import os, subprocess, sys, tempfile, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_py
from gsmfretda_b200 import synth
cfg, bursts, R = int(sys.argv[1]), 20000, 16
w = synth.workload(cfg, n_bursts=bursts, R=R); b = w.bursts()
with tempfile.TemporaryDirectory() as td:
    out = os.path.join(td, "ref.npz")
    subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "refgpu_run.py"), "--cfg", str(cfg), "--bursts", str(bursts),
                    "--repeats", str(R), "--evals", "1", "--out", out], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    g = np.load(out)["mcE"][0]
h = oracle_py.Ref().mc_kernel_sobol(b, w.n_states, w.args[0], R, evals=1)[0]
bad = g != h
t = np.arange(g.size)
print("mismatch frac", bad.mean())
print("by chunk of 20000 threads:", [round(float(bad[i*20000:(i+1)*20000].mean()), 3) for i in range(16)])
print("by repeat:", [round(float(bad[t % R == r].mean()), 3) for r in range(R)])
ph = np.repeat(b.ph_count, R)
for lo, hi in ((0, 30), (30, 50), (50, 80), (80, 150), (150, 1000)):
    m = (ph >= lo) & (ph < hi)
    print(f"ph in [{lo},{hi}): n={int(m.sum())} mismatch {float(bad[m].mean()):.3f}")
d = np.abs(g[bad] - h[bad])
print("abs diff quantiles", np.quantile(d, [0.1, 0.5, 0.9, 0.99]).tolist())
print("first 12 mismatches (t, gpu, host):", [(int(i), float(g[i]), float(h[i])) for i in np.nonzero(bad)[0][:12]])
print("gpu zero frac", float((g == 0).mean()), "host zero frac", float((h == 0).mean()))
print("mean over mismatching: gpu", float(g[bad].mean()), "host", float(h[bad].mean()))
