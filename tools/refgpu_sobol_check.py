#!/usr/bin/env python
"""Does the host build of the reference's sources (oracle/_ref/libref_cpu.so) compute what the
reference's class mc computes on the GPU when it is given cuRAND's own Sobol sequences?
Per-sample comparison of mcE from oracle/refgpu_run.py (GPU, unmodified mc.cu) with
Ref.mc_kernel_sobol (host, same sources, Sobol64 restated from curand_kernel.h).
usage: python tools/refgpu_sobol_check.py [--cfgs 1,2,3,5] [--bursts 20000] [--repeats 16]"""
import argparse
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def compare(cfg, bursts, repeats, evals=2, timeout=200.0):
    import oracle_py
    from gsmfretda_b200 import synth
    from scipy import stats
    w = synth.workload(cfg, n_bursts=bursts, R=repeats)
    b = w.bursts()
    with tempfile.TemporaryDirectory() as td:
        out = os.path.join(td, "ref.npz")
        subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "refgpu_run.py"), "--cfg", str(cfg),
                        "--bursts", str(bursts), "--repeats", str(repeats), "--evals", str(evals),
                        "--out", out], check=True, timeout=timeout, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        d = np.load(out)
        gpu = d["mcE"]
        R = int(d["R"])
    host = oracle_py.Ref().mc_kernel_sobol(b, w.n_states, w.args[0], R, evals=evals)
    res = {"cfg": cfg, "bursts": b.n_bursts, "R": R, "evals": evals}
    for ev in range(evals):
        g, h = gpu[ev], host[ev]
        same = (g == h) | (np.isnan(g) & np.isnan(h))
        close = same | (np.abs(g - h) <= 1e-6)
        ks = stats.ks_2samp(g, h, method="asymp")
        res[f"eval{ev}"] = {"identical_bits_frac": float(same.mean()), "within_1e-6_frac": float(close.mean()),
                            "mean_gpu": float(np.nanmean(g)), "mean_host": float(np.nanmean(h)),
                            "ks_D": float(ks.statistic), "ks_p": float(ks.pvalue)}
    return res


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfgs", default="1,2,3,5")
    ap.add_argument("--bursts", type=int, default=20000)
    ap.add_argument("--repeats", type=int, default=16)
    a = ap.parse_args()
    for c in a.cfgs.split(","):
        print(json.dumps(compare(int(c), a.bursts, a.repeats)), flush=True)
