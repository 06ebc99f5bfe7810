#!/usr/bin/env python
"""Diagnostics for GSM_MODE_NORMAL against the oracle (independent draws): KS on E, two-sample
chi-square homogeneity of the E histograms, spread of the reduced chi-square over epochs/seeds.
A replay-mode run with yet another seed is the control (it is the oracle's own arithmetic).
usage: python tools/normal_diag.py [--cases 1:16000:64,2:10000:100,...]"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
NB, LP = 200, 2


def homogeneity(a, b, min_count=20):
    """two-sample chi-square on bins whose pooled count >= min_count; returns stat, dof, p"""
    from scipy import stats
    a = a[:NB].astype(np.float64); b = b[:NB].astype(np.float64)
    m = (a + b) >= min_count
    a, b = a[m], b[m]
    ka, kb = np.sqrt(b.sum() / a.sum()), np.sqrt(a.sum() / b.sum())
    st = float((((ka * a - kb * b) ** 2) / (a + b)).sum())
    dof = int(m.sum()) - 1
    return st, dof, float(stats.chi2.sf(st, dof))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", default="1:16000:64,2:10000:100,3:60000:16,5:120000:8")
    ap.add_argument("--epochs", type=int, default=6)
    ap.add_argument("--seeds", type=int, default=4)
    a = ap.parse_args()
    import oracle_py
    from scipy import stats
    from gsmfretda_b200 import capi, synth
    O = oracle_py.Oracle()
    for case in a.cases.split(","):
        cfg, nb, R = [int(x) for x in case.split(":")]
        w = synth.workload(cfg, n_bursts=nb, R=R)
        b = w.bursts()
        oh = O.observed_hist(b, NB)
        res = {"cfg": cfg, "bursts": nb, "R": R}
        orc = [O.eval(b, w.n_states, w.args[0], R, 202 + 101 * s, 1) for s in range(a.seeds)]
        ohists = [O.hist(o["mcE"], NB) for o in orc]
        res["chi2_oracle"] = [float(O.chi2(oh, h, NB, R, LP, w.n_states)) for h in ohists]
        for mode, name in ((capi.GSM_MODE_NORMAL, "normal"), (capi.GSM_MODE_REPLAY, "replay")):
            with capi.Mc(0, 1, keep_mce=True, keep_counts=True) as mc:
                mc.set_reSampleTimes(R); mc.set_seed(101); mc.set_hist(NB, LP); mc.set_mode(mode)
                mc.init_bursts(b)
                chis, hists = [], []
                for ep in range(1, a.epochs + 1):
                    c, h = mc.evaluate(w.n_states, w.args[0], sid=0, epoch=ep)
                    chis.append(float(c)); hists.append(h.copy())
                    if ep == 1:
                        e = mc.hmcE(0, nb * R).copy()
                        sa = mc.counts(0, nb * R).copy()
                        ns = mc.steps(0, nb * R).copy()
            k = stats.ks_2samp(e, orc[0]["mcE"], method="asymp")
            res[name] = {
                "chi2": chis, "ks": [float(k.statistic), float(k.pvalue)],
                "homog_vs_oracle0": homogeneity(hists[0], ohists[0]),
                "homog_pooled": homogeneity(np.sum(hists, axis=0), np.sum(ohists, axis=0)),
                "mean_E": [float(e.mean()), float(orc[0]["mcE"].mean())],
                "mean_sum_ai": [float(sa.mean()), float(orc[0]["sum_ai"].mean())],
                "mean_steps": [float(ns.mean()), float(orc[0]["n_steps"].mean())],
            }
        res["homog_oracle0_vs_oracle1"] = homogeneity(ohists[0], ohists[1])
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
