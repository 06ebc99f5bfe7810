#!/usr/bin/env python
"""Record what the REFERENCE'S OWN class mc (unmodified mc.cu, cuRAND Sobol64, libdevice;
oracle/_ref/libref_gpu.so) computes on a B200 for small synthetic workloads:
    gpurun -- python tools/make_refgpu_golden.py        # writes gpurun_out/refgpu_golden/*.npz
The files are committed under tests/golden/refgpu/ and pin the host build of the same sources
(tests/test_reference_gpu_golden.py): the reference run here, on its own generator."""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out", "refgpu_golden")
CASES = [(1, 600, 4), (2, 600, 4), (3, 400, 16), (5, 400, 4), (2, 6000, 4)]   # the last one crosses thread 20000 (quirk Q8); R in {4, 16}: the reference did not come back for R = 8 and 64 here

if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    for cfg, nb, R in CASES:
        with tempfile.TemporaryDirectory() as td:
            out = os.path.join(td, "ref.npz")
            subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "refgpu_run.py"), "--cfg", str(cfg), "--bursts", str(nb),
                            "--repeats", str(R), "--evals", "2", "--out", out], check=True, timeout=60)
            d = np.load(out)
            np.savez_compressed(os.path.join(OUT, f"refgpu_cfg{cfg}_{nb}x{R}.npz"), cfg=cfg, n_bursts=nb, R=int(d["R"]),
                                mcE=d["mcE"].astype(np.float32))
        print("wrote", cfg, nb, R, flush=True)
