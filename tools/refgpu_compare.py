#!/usr/bin/env python
"""Statistical comparison of this library with the reference's own GPU `class mc`
(oracle/_ref/libref_gpu.so, run in a process of its own by oracle/refgpu_run.py).
usage: python tools/refgpu_compare.py [--cfgs 1,2,3,5] [--bursts 20000] [--repeats 64]
Prints one JSON line per config: KS D/p (E samples), chi-square of both against the observed
histogram, and the same numbers between two evaluations of each implementation (noise floor)."""
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
sys.path.insert(0, os.path.join(ROOT, "tools"))


def ks2(x, y):
    from scipy import stats
    r = stats.ks_2samp(x, y, method="asymp")
    return float(r.statistic), float(r.pvalue)


def hist200(e):
    z = e.astype(np.float64)
    idx = (z[(z >= 0) & (z < 1)] * 200).astype(np.int64)
    return np.bincount(idx, minlength=200).astype(np.uint64)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfgs", default="1,2,3,5")
    ap.add_argument("--bursts", type=int, default=20000)
    ap.add_argument("--repeats", type=int, default=64)
    ap.add_argument("--mode", type=int, default=0, help="gsm mode (0 replay, 1 normal)")
    ap.add_argument("--timeout", type=float, default=150.0, help="seconds allowed for one reference run")
    a = ap.parse_args()
    import oracle_py
    from gsmfretda_b200 import capi, synth
    O = oracle_py.Oracle()
    for cfg in [int(c) for c in a.cfgs.split(",")]:
        w = synth.workload(cfg, n_bursts=a.bursts, R=a.repeats)
        b = w.bursts()
        with tempfile.TemporaryDirectory() as td:
            out = os.path.join(td, "ref.npz")
            subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "refgpu_run.py"), "--cfg", str(cfg),
                            "--bursts", str(a.bursts), "--repeats", str(a.repeats), "--evals", "2",
                            "--out", out], check=True, timeout=a.timeout)
            d = np.load(out)
            ref = d["mcE"]
            R = int(d["R"])
        with capi.Mc(0, 1, keep_mce=True) as mc:
            mc.set_reSampleTimes(R)
            mc.set_seed(777)
            mc.set_hist(200, 2)
            if a.mode and hasattr(mc, "set_mode"):
                mc.set_mode(a.mode)
            mc.init_bursts(b)
            ours = []
            for ep in (1, 2):
                mc.evaluate(w.n_states, w.args[0], sid=0, epoch=ep)
                ours.append(mc.hmcE(0, b.n_bursts * R).copy())
        oh = O.observed_hist(b, 200)
        chi = lambda e: O.chi2(oh, np.concatenate([hist200(e), np.zeros(2, np.uint64)]), 200, R, 2, w.n_states)
        res = {"cfg": cfg, "bursts": b.n_bursts, "R": R, "mode": a.mode,
               "ks_ours_vs_ref": ks2(ours[0], ref[0]), "ks_ours_vs_ours": ks2(ours[0], ours[1]),
               "ks_ref_vs_ref": ks2(ref[0], ref[1]),
               "chi2_ours": [chi(ours[0]), chi(ours[1])], "chi2_ref": [chi(ref[0]), chi(ref[1])],
               "mean_ours": float(ours[0].mean()), "mean_ref": float(ref[0].mean()),
               "nan_ref": int(np.isnan(ref[0]).sum())}
        from normal_diag import homogeneity
        res["homog_ours_vs_ref"] = homogeneity(hist200(ours[0]), hist200(ref[0]))
        res["homog_ref_vs_ref"] = homogeneity(hist200(ref[0]), hist200(ref[1]))
        res["homog_ours_vs_ours"] = homogeneity(hist200(ours[0]), hist200(ours[1]))
        res["chi2_rel_diff"] = abs(res["chi2_ours"][0] - res["chi2_ref"][0]) / res["chi2_ref"][0]
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
