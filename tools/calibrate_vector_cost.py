#!/usr/bin/env python
"""Time single vectors of the config-4 population on the GPU against the cost model behind
sharding.deal_population (vector_cost): prints model cost, stationary exit flux x mean burst
duration, measured ms per evaluation and the least-squares fit  ms = c0 + c1 * flux*T.
usage (GPU box): python tools/calibrate_vector_cost.py [n_vectors]"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from gsmfretda_b200 import capi, synth, sharding

nv = int(sys.argv[1]) if len(sys.argv) > 1 else 24
w = synth.workload(4)
b = w.bursts()
cost = np.array([sharding.vector_cost(2, a) for a in w.args])
order = np.argsort(cost)
pick = order[np.linspace(0, len(order) - 1, nv).astype(int)]
mean_T = float(np.mean(b.T))
mc = capi.Mc(0, 1)
mc.set_reSampleTimes(w.R); mc.set_seed(1); mc.set_hist(200, 0)
mc.init_bursts(b)
rows = []
for i in pick:
    a = w.args[i]
    mc.evaluate(2, a, sid=0)
    ms = []
    for _ in range(5):
        mc.run_resident(b.n_bursts, 0); mc.sync(0); ms.append(mc.elapsed_ms(0))
    k12, k21 = float(a[2]), float(a[3])
    flux = 2.0 * k12 * k21 / (k12 + k21)
    rows.append((int(i), cost[i], flux * mean_T, float(np.median(ms))))
    print("vector %4d model %.1f fluxT %.3f ms %.4f" % rows[-1])
mc.close()
r = np.array(rows)
A = np.stack([np.ones(len(r)), r[:, 2]], 1)
c, *_ = np.linalg.lstsq(A, r[:, 3], rcond=None)
pred = A @ c
print("mean burst duration %.4g s; fit ms = %.4f + %.4f * fluxT; max rel. residual %.3f" % (mean_T, c[0], c[1], float(np.max(np.abs(pred - r[:, 3]) / r[:, 3]))))
m = r[:, 1] / r[:, 1].mean() * r[:, 3].mean()
print("current model (scaled to the same mean): max rel. residual %.3f" % float(np.max(np.abs(m - r[:, 3]) / r[:, 3])))
