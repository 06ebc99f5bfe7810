#!/usr/bin/env python
"""Parameter evaluations per second at the reference's own scale (config 1: 7 733 bursts x 4
repeats) with S stream slots in flight, as gpuWorker drives them: one host thread issues
set_params + run_kernel round-robin and polls streamQuery.  usage: python tools/multistream_probe.py"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gsmfretda_b200 import capi, synth
w = synth.workload(1); b = w.bursts()
for mode, nm in ((capi.GSM_MODE_REPLAY, "replay"), (capi.GSM_MODE_NORMAL, "normal")):
    for S in (1, 2, 4, 8, 16):
        with capi.Mc(0, S) as mc:
            mc.set_reSampleTimes(w.R); mc.set_hist(200, 2); mc.set_mode(mode); mc.init_bursts(b)
            for sid in range(S):
                mc.set_nstates(w.n_states, sid); mc.setBurstBd(0, -1, sid)
            total = 4000
            issued = done = 0
            busy = [False] * S
            for sid in range(S):   # warm
                mc.set_params(w.n_states, sid, w.args[0]); mc.run_kernel(b.n_bursts, sid)
            for sid in range(S):
                mc.sync(sid)
            t0 = time.perf_counter()
            while done < total:
                for sid in range(S):
                    if busy[sid] and mc.streamQuery(sid):
                        mc.chi2(sid); busy[sid] = False; done += 1
                    if not busy[sid] and issued < total:
                        mc.set_params(w.n_states, sid, w.args[0]); mc.run_kernel(b.n_bursts, sid)
                        busy[sid] = True; issued += 1
            dt = time.perf_counter() - t0
            print(f"{nm} streams {S:2d}: {total / dt:9.0f} param-evals/s  ({dt / total * 1e6:.1f} us per evaluation, {total * b.n_bursts * w.R / dt / 1e9:.2f} G burst-samples/s)", flush=True)
