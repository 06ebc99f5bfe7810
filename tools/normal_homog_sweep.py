#!/usr/bin/env python
"""Histogram homogeneity (chi-square, p) of normal mode vs replay mode over several seeds, with
replay-vs-replay as the control: is a small p of the normal-mode test noise or bias?
usage: python tools/normal_homog_sweep.py [cfg bursts R]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from gsmfretda_b200 import capi, synth
from test_gpu_normal_mode import homogeneity
cfg, nb, R = (int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (3, 60000, 16)))
w = synth.workload(cfg, n_bursts=nb, R=R); b = w.bursts()
def hists(mode, seeds):
    out = []
    with capi.Mc(0, 1) as mc:
        mc.set_reSampleTimes(R); mc.set_hist(200, 2); mc.set_mode(mode); mc.init_bursts(b)
        for s in seeds:
            mc.set_seed(s); out.append(mc.evaluate(w.n_states, w.args[0], epoch=1)[1].copy())
    return out
K = 12
hn = hists(capi.GSM_MODE_NORMAL, range(100, 100 + K))
hr = hists(capi.GSM_MODE_REPLAY, range(500, 500 + K))
hr2 = hists(capi.GSM_MODE_REPLAY, range(900, 900 + K))
pn = [homogeneity(a, c)[1] for a, c in zip(hn, hr)]
pc = [homogeneity(a, c)[1] for a, c in zip(hr2, hr)]
print("normal vs replay p:", np.round(sorted(pn), 4).tolist())
print("replay vs replay p:", np.round(sorted(pc), 4).tolist())
tot_n, tot_r, tot_r2 = sum(hn), sum(hr), sum(hr2)
print("pooled (12x samples) normal vs replay:", homogeneity(tot_n, tot_r), " replay vs replay:", homogeneity(tot_r2, tot_r))
