#!/usr/bin/env python
"""oracle/refgpu_run.py — TEST INFRASTRUCTURE.  Runs the REFERENCE'S OWN `class mc` (unmodified
mc.cu built for sm_100a, oracle/_ref/libref_gpu.so) on a synthetic workload in a process of
its own (its destructor resets the device) and writes what the checks need:

    python oracle/refgpu_run.py --cfg 2 --bursts 20000 --repeats 64 --evals 3 --out x.npz [--time K]

  x.npz: mcE[evals, N*R] float32 (the reference's hmcE after each evaluation; its Sobol
         states persist from one evaluation to the next, quirk Q8), R, N
  --time K additionally prints one JSON line with the reference's burst-samples/s through its own
         API (set_params ... run_kernel + D2H + streamQuery; the host histogram + chi-square of
         streamWorker.cpp:187-206 is timed too, through the reference's vendored boost::histogram)
"""
import argparse
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfg", type=int, default=2)
    ap.add_argument("--bursts", type=int, default=20000)
    ap.add_argument("--repeats", type=int, default=64)
    ap.add_argument("--evals", type=int, default=1)
    ap.add_argument("--pvec", type=int, default=0, help="which parameter vector of the workload")
    ap.add_argument("--out", default=None)
    ap.add_argument("--time", type=int, default=0)
    a = ap.parse_args()

    import oracle_py
    from gsmfretda_b200 import synth

    STAGES = {}
    w = synth.workload(a.cfg, n_bursts=a.bursts, R=a.repeats)
    b = w.bursts()
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(1)
    os.dup2(devnull, 1)          # the reference prints device info and "blockSize" on stdout
    try:
        def stage(msg, t=[time.perf_counter()]):
            now = time.perf_counter()
            print("[refgpu_run] %s (+%.2f s)" % (msg, now - t[0]), file=sys.stderr, flush=True)
            STAGES[msg] = round(now - t[0], 3)
            t[0] = now
        m = oracle_py.RefGpu(0, 1)
        stage("created")
        R = m.set_reSampleTimes(a.repeats)
        m.init_bursts(b)
        stage("init_data_gpu done, R=%d" % R)
        arg = w.args[a.pvec]
        out = []
        for i in range(a.evals):
            out.append(m.evaluate(w.n_states, arg))
            stage("evaluation %d done" % i)
        line = None
        if a.time > 0:
            m.time(w.n_states, arg, 3)
            s = m.time(w.n_states, arg, a.time)
            # the host half the reference does per evaluation (streamWorker.cpp:187-206): its own
            # boost::histogram (the vendored copy) + chi-square loop, compiled into libref_gpu.so
            host_total, chi2 = m.host_half(b.n_bursts, w.n_states, 3)
            host_s = host_total / 3
            line = {"impl": "reference class mc on the GPU (oracle/_ref/libref_gpu.so)", "bursts": b.n_bursts,
                    "repeats": R, "evals_timed": a.time, "kernel_plus_d2h_s_per_eval": s / a.time,
                    "host_hist_s_per_eval": host_s, "chi2": chi2, "stages_s": STAGES,
                    "burst_samples_per_s_kernel_d2h": b.n_bursts * R * a.time / s,
                    "burst_samples_per_s_with_host_hist": b.n_bursts * R / (s / a.time + host_s)}
    finally:
        sys.stdout.flush()
        os.dup2(saved, 1)
    if a.out:
        np.savez_compressed(a.out, mcE=np.stack(out), R=R, N=b.n_bursts)
    if line:
        print(json.dumps(line), flush=True)
    os._exit(0)   # skip the reference's destructor path (cudaDeviceReset) and ctypes teardown


if __name__ == "__main__":
    main()
