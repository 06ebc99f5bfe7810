"""The reference's OWN GPU output pins the checker.

tests/golden/refgpu/*.npz hold what the reference's unmodified `class mc` (mc.cu + cuRAND
scrambled Sobol64 + libdevice, built for sm_100a as oracle/_ref/libref_gpu.so) computed on a B200
(tools/make_refgpu_golden.py).  The host build of the same sources (oracle/_ref/libref_cpu.so)
must reproduce those per-sample efficiencies BIT FOR BIT when its cuRAND stand-in replays cuRAND's
own Sobol64 sequences (restated from the toolkit's curand_kernel.h; tables from libcurand's host
API) — both evaluations, i.e. including the RNG state the reference carries from one evaluation to
the next (quirk Q8), thread ids beyond the 20000 Sobol dimensions, and the background path.

Chain of custody of a parity claim:
    reference on the GPU  ==  reference sources on the host (Sobol draws)        [this file]
    reference sources on the host (contract draws)  ==  C restatement            [test_oracle_vs_reference.py]
    C restatement  ==  CUDA kernel through the C ABI                             [test_gpu_parity.py]
so the only thing that separates this library from the reference is the source of the uniform
words (counter-based Philox instead of 1-D Sobol sequences).
"""
import glob
import os

import numpy as np
import pytest

from gsmfretda_b200 import synth

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "refgpu", "*.npz")))


def test_refgpu_golden_present():
    assert len(GOLD) >= 5


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p)[:-4] for p in GOLD])
def test_host_build_replays_the_reference_gpu_bit_for_bit(ref, path):
    d = np.load(path)
    cfg, nb, R = int(d["cfg"]), int(d["n_bursts"]), int(d["R"])
    w = synth.workload(cfg, n_bursts=nb, R=R)
    b = w.bursts()
    want = d["mcE"]
    got = ref.mc_kernel_sobol(b, w.n_states, w.args[0], R, evals=want.shape[0])
    assert got.shape == want.shape
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), \
        f"{(got != want).mean():.4%} of the samples differ from the reference's GPU run"
    assert not np.array_equal(want[0], want[1])      # the second evaluation continues the sequences


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p)[:-4] for p in GOLD])
def test_c_restatement_replays_the_reference_gpu_bit_for_bit(oracle, path):
    """The oracle itself (oracle/pda_oracle.c, no reference source involved) on the reference's
    generators: equal to the reference's GPU run, so the checker is pinned on any box."""
    d = np.load(path)
    cfg, nb, R = int(d["cfg"]), int(d["n_bursts"]), int(d["R"])
    w = synth.workload(cfg, n_bursts=nb, R=R)
    want = d["mcE"]
    got = oracle.eval_sobol(w.bursts(), w.n_states, w.args[0], R, evals=want.shape[0])
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), f"{(got != want).mean():.4%} differ"


@pytest.mark.parametrize("cfg", [1, 2, 3, 5])
def test_c_restatement_equals_host_build_on_sobol(oracle, ref, cfg):
    w = synth.workload(cfg, n_bursts=6000, R=8)      # 48 000 threads: beyond the 20 000 dimensions
    b = w.bursts()
    a = oracle.eval_sobol(b, w.n_states, w.args[0], 8, evals=3)
    r = ref.mc_kernel_sobol(b, w.n_states, w.args[0], 8, evals=3)
    assert np.array_equal(a.view(np.uint32), r.view(np.uint32))
    h = oracle.eval_sobol(b, w.n_states, w.args[0], 8, evals=1, N=2500)   # N only sets the thread count (Q7)
    assert np.array_equal(h, ref.mc_kernel_sobol(b, w.n_states, w.args[0], 8, evals=1, N=2500))


def test_sobol_tables_are_curands():
    """First words of dimension 0/1 after curand_init(..., offset 1234), read off a B200
    (tools/sobol_probe.cu)."""
    import oracle_py
    dv, sc = oracle_py.sobol_tables()

    def first(idx):
        i, x = 1234, int(sc[idx])
        g = i ^ (i >> 1)
        for k in range(64):
            if g >> k & 1:
                x ^= int(dv[idx, k])
        return x
    assert first(0) == 0x4d14c2df07cb3028 and first(1) == 0x8e2c652faa03ff04
    assert first(3) == 0x2855b4fb409ff155 and first(12) == 0x73462c7a37d15205


@pytest.mark.parametrize("name", ["edge_mixed_p0", "edge_far_ticks_p0"])
def test_sobol_draw_source_on_edge_bursts(oracle, ref, name):
    """Empty, zero-duration and one-photon bursts, ticks far from 0: a thread whose burst has no
    duration still takes the initial-state draw (mc.cu:110 precedes the loop)."""
    from helpers import load_golden
    b, d = load_golden(name)
    n, R = int(d["n"]), int(d["R"])
    a = oracle.eval_sobol(b, n, d["args"], R, evals=3)
    r = ref.mc_kernel_sobol(b, n, d["args"], R, evals=3)
    assert np.array_equal(a.view(np.uint32), r.view(np.uint32))
