#!/usr/bin/env python
"""Where a small evaluation (the reference's own scale: a few thousand bursts x 4 repeats) spends
its time: device time per evaluation against the number of bursts, warm and cold L2.
usage: python tools/small_job_probe.py [cfg]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gsmfretda_b200 import capi, synth  # noqa: E402

cfg = int(sys.argv[1]) if len(sys.argv) > 1 else 1
w = synth.workload(cfg)
b = w.bursts()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for mode in (capi.GSM_MODE_REPLAY, capi.GSM_MODE_NORMAL):
    with capi.Mc(0, 2) as mc:
        mc.set_reSampleTimes(w.R); mc.set_seed(1); mc.set_hist(200, 2); mc.set_mode(mode)
        mc.init_bursts(b)
        mc.evaluate(w.n_states, w.args[0])
        for N in (0, 8, 100, 1000, b.n_bursts):
            for cold in (False, True):
                ms = []
                for i in range(60):
                    if cold:
                        flush.zero_(); torch.cuda.synchronize()
                    mc.run_resident(N, 0); mc.sync(0)
                    ms.append(mc.elapsed_ms(0))
                ms = np.array(ms[10:]) * 1e3
                print(f"mode {mode} N {N:6d} R {w.R} {'cold' if cold else 'warm'}: median {np.median(ms):7.1f} us  min {ms.min():7.1f} us")
        # host-side cost of one evaluation through the public API (warm)
        t0 = time.perf_counter()
        for i in range(500):
            mc.evaluate(w.n_states, w.args[0])
        print(f"mode {mode} public API evaluate(): {(time.perf_counter() - t0) / 500 * 1e6:.1f} us per evaluation (warm L2)")
