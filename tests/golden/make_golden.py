#!/usr/bin/env python
"""Generate the golden vectors of tests/golden/ from the REFERENCE'S OWN code.

Runs only in the build container, where /root/reference exists: oracle/Makefile compiles the
reference's mc_kernel / gen_rand.cuh / binom.cuh / eigenhelper.cpp for the host
(oracle/_ref/libref_cpu.so, cuRAND bound to the draw contract by oracle/ref_shim.cpp), and this
script records its inputs and outputs.  The reference's tests hold no golden vectors for this
path, so these are "outputs of the reference itself run here".

    python tests/golden/make_golden.py        # rewrites tests/golden/*.npz

Each case stores the burst arrays, the parameter vector, (R, seed, epoch) and, from the
reference: E, v, K, P_i2j, p of set_params and the per-sample mcE of mc_kernel.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import oracle_py  # noqa: E402
from gsmfretda_b200 import synth  # noqa: E402

CASES = [
    # name, cfg, bursts, R (power of two: mc.cu:101), seed, epoch
    ("cfg1_static2", 1, 400, 4, 101, 0),
    ("cfg2_dynamic2", 2, 300, 8, 202, 1),
    ("cfg3_fullK3", 3, 300, 4, 303, 2),
    ("cfg4_population2", 4, 200, 4, 404, 3),
    ("cfg5_bg4", 5, 300, 4, 505, 4),
]


def special_bursts():
    """Edge cases: empty bursts, zero duration, a burst starting at tick 0, one-photon bursts,
    long acceptor runs (BTPE: n*p > 30), ticks far from zero (float rounding of the edges)."""
    b = synth.make_bursts(64, (0.05, 0.95), seed=77, first_start=0, max_extra=600, mean_extra=300.0)
    ph = b.ph_count.copy()
    ph[3] = 0
    ph[7] = 1
    ph[11] = 2
    T = b.T.copy()
    T[5] = 0.0
    b.ph_count = ph
    b.T = T
    far = synth.make_bursts(32, (0.5,), seed=78, first_start=3_000_000_000_000)
    return b, far


def main():
    R = oracle_py.Ref()
    for name, cfg, nb, rep, seed, epoch in CASES:
        w = synth.workload(cfg, n_bursts=nb, R=rep, P=3 if cfg == 4 else None)
        b = w.bursts()
        for pi in range(w.P):
            args = w.args[pi]
            ok, E, v, K, Pij, p = R.params(w.n_states, args)
            assert ok
            mcE = R.mc_kernel(b, w.n_states, args, rep, seed, epoch + pi)
            np.savez_compressed(
                os.path.join(HERE, f"{name}_p{pi}.npz"), start=b.start, stop=b.stop, istart=b.istart,
                ph_count=b.ph_count, T=b.T, burst_ad=b.burst_ad, burst_dd=b.burst_dd, SgDivSr=b.SgDivSr,
                clk_p=b.clk_p, bg_ad_rate=b.bg_ad_rate, bg_dd_rate=b.bg_dd_rate, n=w.n_states, args=args,
                R=rep, seed=seed, epoch=epoch + pi, ref_E=E, ref_v=v, ref_K=K, ref_P_i2j=Pij, ref_p=p,
                ref_mcE=mcE)
    sp, far = special_bursts()
    for name, b, n, args in (
            ("edge_mixed", sp, 2, synth.pack_args([0.5, 0.45], [40, 60], [1, 1])),
            ("edge_far_ticks", far, 3, synth.pack_args([0.2, 0.5, 0.8], [900, 2500, 1800, 4000, 3200, 1400], [2, 2, 2]))):
        ok, E, v, K, Pij, p = R.params(n, args)
        mcE = R.mc_kernel(b, n, args, 8, 909, 9)
        np.savez_compressed(
            os.path.join(HERE, f"{name}_p0.npz"), start=b.start, stop=b.stop, istart=b.istart,
            ph_count=b.ph_count, T=b.T, burst_ad=b.burst_ad, burst_dd=b.burst_dd, SgDivSr=b.SgDivSr,
            clk_p=b.clk_p, bg_ad_rate=b.bg_ad_rate, bg_dd_rate=b.bg_dd_rate, n=n, args=args, R=8,
            seed=909, epoch=9, ref_E=E, ref_v=v, ref_K=K, ref_P_i2j=Pij, ref_p=p, ref_mcE=mcE)
    # n = 5, 6 (the top of the reference's own test matrix, tests/gen_rand_tests.cpp:79-82) and
    # n = 8 (the top of get_max_conversion_capacity's default range), full K matrices
    for n, bg, seed in ((5, 0.0, 515), (6, 1.0, 616), (8, 0.0, 818)):
        rng = np.random.default_rng(9000 + n)
        E = np.linspace(0.08, 0.92, n)
        k = np.exp(rng.uniform(np.log(150.0), np.log(4000.0), size=n * n - n))
        args = synth.pack_args(E, k, rng.uniform(1.0, 3.0, size=n))
        b = synth.make_bursts(160, E, seed=seed, bg_ad_rate=1200.0 * bg, bg_dd_rate=2000.0 * bg)
        ok, Ee, v, K, Pij, pp = R.params(n, args)
        assert ok
        mcE = R.mc_kernel(b, n, args, 4, seed, n)
        np.savez_compressed(
            os.path.join(HERE, f"n{n}_fullK{'_bg' if bg else ''}_p0.npz"), start=b.start, stop=b.stop, istart=b.istart,
            ph_count=b.ph_count, T=b.T, burst_ad=b.burst_ad, burst_dd=b.burst_dd, SgDivSr=b.SgDivSr,
            clk_p=b.clk_p, bg_ad_rate=b.bg_ad_rate, bg_dd_rate=b.bg_dd_rate, n=n, args=args, R=4,
            seed=seed, epoch=n, ref_E=Ee, ref_v=v, ref_K=K, ref_P_i2j=Pij, ref_p=pp, ref_mcE=mcE)
    print("golden vectors written to", HERE)


if __name__ == "__main__":
    main()
