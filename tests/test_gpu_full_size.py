"""The two BASELINE.json configurations whose full size had no oracle check so far, launched exactly
as bench.py launches them and compared with the oracle bit for bit:

  * config 5 at 1e7 bursts x 4 repeats (the fixed set of 8 blocks of 1.25e6 bursts bench.py times):
    whole histogram and chi-square bits;
  * config 4 as ONE gsm_run_batch of 1024 vectors over 1e5 bursts x 16 repeats — the launch shape with
    span tickets, warps changing vector and per-vector epochs: the first and last vectors, the ones
    either side of every boundary gsm_launch_shape reports, and a spread in between.
"""
import ctypes as C
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

from gsmfretda_b200 import synth
from helpers import NB, LEFT_PAD

pytestmark = pytest.mark.gpu

CORES = len(os.sched_getaffinity(0))


def test_config5_at_ten_million_bursts_equals_the_oracle(gsm, oracle):
    w = synth.workload(5)
    blocks, each = 8, w.n_bursts // 8
    assert w.n_bursts == 10_000_000 and w.R == 4
    with ThreadPoolExecutor(4) as ex:           # block j: seed + 1000 j, as bench.gen_blocks makes the set
        parts = list(ex.map(lambda j: w.bursts(n_bursts=each, seed=w.seed + 1000 * j), range(blocks)))
    b = synth.concat(parts)
    del parts
    assert b.n_bursts == 10_000_000
    with gsm.Mc(0, 1) as mc:
        mc.set_reSampleTimes(w.R); mc.set_seed(77); mc.set_hist(NB, LEFT_PAD)
        mc.init_bursts(b)
        assert mc.layout() == 2
        c, h = mc.evaluate(w.n_states, w.args[0], sid=0, epoch=9)
        assert mc.flags(0) == 0
    assert int(h.sum()) == b.n_bursts * w.R
    co, ho = oracle.eval_chi2(b, w.n_states, w.args[0], w.R, 77, 9, oracle.observed_hist(b, NB), NB, LEFT_PAD,
                              nthreads=CORES)
    assert np.array_equal(h, ho), int((h != ho).sum())
    assert np.float32(c).view(np.uint32) == np.float32(co).view(np.uint32)


def test_config4_batch_of_1024_vectors_equals_the_oracle(gsm, oracle):
    w = synth.workload(4)
    b = w.bursts()
    B, R, n, P = b.n_bursts, w.R, w.n_states, w.args.shape[0]
    assert (B, R, P) == (100_000, 16, 1024)
    # where the tickets of this launch fall: a ticket is `span` consecutive chunks of one vector
    shape = (C.c_int64 * 6)()
    assert gsm.load().gsm_launch_shape(B, R, P, 148 * 4, 0, shape) == 0
    chunk, span, _, _, per_p, _ = [int(v) for v in shape]
    tickets_per_vector = -(-per_p // span)
    # vectors whose first ticket is handed out right after a resident wave of the persistent grid ended
    # (148 SMs x 4 blocks x 8 warps tickets in flight), and their neighbours
    wave = 148 * 4 * 8
    at_waves = sorted({min(P - 1, (k * wave) // tickets_per_vector + d) for k in range(1, 4) for d in (0, 1)})
    picks = sorted(set([0, 1, 2, P // 2 - 1, P // 2, P - 3, P - 2, P - 1] + at_waves +
                       [int(x) for x in np.random.default_rng(4).integers(3, P - 3, 6)]))
    assert len(picks) >= 16
    epoch0 = 40
    with gsm.Mc(0, 2) as mc:
        mc.set_reSampleTimes(R); mc.set_seed(77); mc.set_hist(NB, LEFT_PAD)
        mc.init_bursts(b)
        mc.set_epoch(0, epoch0)
        mc.run_batch(B, 0, n, w.args)                   # what bench.py's population leg launches
        mc.sync(0)
        assert mc.flags(0) == 0
        chi2 = mc.chi2v(0).copy()
        hists = {p: mc.hist(0, p).copy() for p in picks}
        mass = [int(mc.hist(0, p).sum()) for p in range(P)]
    assert chi2.shape == (P,) and mass == [B * R] * P   # every vector got every sample exactly once
    ohist = oracle.observed_hist(b, NB)
    for p in picks:                                     # vector p draws with epoch + p
        co, ho = oracle.eval_chi2(b, n, w.args[p], R, 77, epoch0 + p, ohist, NB, LEFT_PAD, nthreads=CORES)
        assert np.array_equal(hists[p], ho), (p, int((hists[p] != ho).sum()), chunk, span)
        assert np.float32(chi2[p]).view(np.uint32) == np.float32(co).view(np.uint32), p
