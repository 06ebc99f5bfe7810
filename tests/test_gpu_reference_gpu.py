"""Live on the GPU box: the reference's own `class mc` (oracle/_ref/libref_gpu.so, a process of
its own) against the host build of the same sources on cuRAND's Sobol sequences, per sample, at
sizes beyond the committed fixtures; and the throughput of the two GPU implementations side by side."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("cfg", [1, 2, 3, 5])
def test_reference_gpu_equals_host_build_on_sobol(ref, cfg):
    import oracle_py
    if not oracle_py.have_refgpu():
        pytest.skip("oracle/_ref/libref_gpu.so not built")
    from refgpu_sobol_check import compare
    r = compare(cfg, 20000, 16, evals=2)
    for ev in ("eval0", "eval1"):
        # libdevice vs explicit-rounding elementary functions: at most a handful of 320 000 samples
        assert r[ev]["identical_bits_frac"] >= 1.0 - 2e-5, r
