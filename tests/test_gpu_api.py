"""GPU: device primitives against the oracle bit for bit, the batch / resident / multi-stream
entry points, edge cases, error behaviour, and full-size property checks."""
import ctypes as C

import os

import numpy as np
import pytest

from gsmfretda_b200 import synth
from helpers import NB, LEFT_PAD, bits

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mc1(gsm):
    m = gsm.Mc(0, 1)
    yield m
    m.close()


def test_probe_philox_and_math_bit_exact(mc1, oracle):
    rng = np.random.default_rng(1)
    n = 4096
    ctr = rng.integers(0, 2 ** 32, 4 * n, dtype=np.uint64).astype(np.uint32)
    key = rng.integers(0, 2 ** 32, 2 * n, dtype=np.uint64).astype(np.uint32)
    got = mc1.probe(0, n, w0=ctr, w1=key, out="ou", out_len=4 * n)
    L = oracle.lib
    want = np.zeros(4 * n, np.uint32)
    for i in range(n):
        c = (C.c_uint32 * 4)(*[int(v) for v in ctr[4 * i:4 * i + 4]]); k = (C.c_uint32 * 2)(*[int(v) for v in key[2 * i:2 * i + 2]])
        o = (C.c_uint32 * 4)()
        L.pda_oracle_philox(c, k, o)
        want[4 * i:4 * i + 4] = list(o)
    assert np.array_equal(got, want)
    kat = mc1.probe(0, 1, w0=np.zeros(4, np.uint32), w1=np.zeros(2, np.uint32), out="ou", out_len=4)
    assert [int(v) for v in kat] == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]

    w = rng.integers(0, 2 ** 32, n, dtype=np.uint64).astype(np.uint32)
    w[:6] = [0, 1, 0x7fffffff, 0x80000000, 0x80000001, 0xffffffff]
    assert np.array_equal(bits(mc1.probe(6, n, w0=w, out="of")), bits([L.pda_oracle_u01(int(v)) for v in w]))
    assert np.array_equal(bits(mc1.probe(4, n, w0=w, out="of")), bits([L.pda_oracle_norm_icdf(int(v)) for v in w]))
    x = (rng.random(n) ** 3 * 10 + 1e-12).astype(np.float32)
    assert np.array_equal(bits(mc1.probe(1, n, f0=x, out="of")), bits([L.pda_oracle_logf(float(v)) for v in x]))
    assert np.array_equal(bits(mc1.probe(5, n, f0=x, out="of")), bits([L.pda_oracle_root6f(float(v)) for v in x]))
    d = rng.random(n) ** 4 + 1e-300
    assert np.array_equal(mc1.probe(2, n, d0=d, out="od").view(np.uint64),
                          np.array([L.pda_oracle_log(float(v)) for v in d]).view(np.uint64))
    e = rng.uniform(-700, 700, n)
    e[:3] = [0.0, -0.0, -500.0]
    assert np.array_equal(mc1.probe(3, n, d0=e, out="od").view(np.uint64),
                          np.array([L.pda_oracle_exp(float(v)) for v in e]).view(np.uint64))
    k = (rng.random(n) * 5000 + 1e-3).astype(np.float32)
    assert np.array_equal(bits(mc1.probe(10, n, w0=w, f0=k, out="of")),
                          bits([L.pda_oracle_draw_tau(float(a), int(b)) for a, b in zip(k, w)]))
    E = rng.uniform(0.02, 0.98, n).astype(np.float32); v = rng.uniform(0, 6, n).astype(np.float32)
    assert np.array_equal(bits(mc1.probe(9, n, w0=w, f0=E, f1=v, out="of")),
                          bits([L.pda_oracle_draw_e(float(a), 52.0, float(b), int(c)) for a, b, c in zip(E, v, w)]))
    hx = np.concatenate([rng.uniform(-0.2, 1.2, n - 4), [1.0, np.nan, 0.0, -0.0]]).astype(np.float32)
    hi = mc1.probe(13, n, ni=NB, f0=hx, out="oi")
    want_h = np.zeros(NB + 2, np.uint64)
    np.add.at(want_h, hi, 1)
    assert np.array_equal(want_h, oracle.hist(hx, NB))


def test_probe_binomial_poisson_bit_exact(mc1, oracle):
    rng = np.random.default_rng(2)
    L = oracle.lib
    n = 3000
    seeds = rng.integers(1, 2 ** 63, n).astype(np.uint64)
    w = np.stack([(seeds & 0xffffffff).astype(np.uint32), (seeds >> 32).astype(np.uint32)], 1).reshape(-1)
    nn = rng.choice([-3, -1, 0, 1, 2, 3, 5, 9, 20, 70, 150, 400, 3000], n).astype(np.int32)
    pp = rng.choice([1e-3, 0.05, 0.3, 0.5, 0.62, 0.9, 0.999], n)
    got = mc1.probe(7, n, ni=4, w0=w, i0=nn, d0=pp, out="ol")
    st = (C.c_uint32 * 4)()
    want = np.zeros(n, np.int64)
    for i in range(n):
        L.pda_oracle_rk_seed(int(seeds[i]), st)
        s = 0
        for _ in range(4):
            s = (s * 31 + L.pda_oracle_binomial(st, int(nn[i]), float(pp[i])))
            s = (s + 2 ** 63) % 2 ** 64 - 2 ** 63          # int64 wrap like the device
        want[i] = s
    assert np.array_equal(got, want)
    lam = rng.choice([1e-4, 0.02, 0.7, 3.0, 40.0, 480.0, 510.0, 990.0], n)
    got = mc1.probe(8, n, w0=w, d0=lam, out="ou")
    want = np.array([L.pda_oracle_poisson(int(w[2 * i]), int(w[2 * i + 1]), float(lam[i])) for i in range(n)], np.uint32)
    assert np.array_equal(got, want)


def test_probe_inversion_division_is_the_ieee_quotient(mc1):
    """div_pos_normal (the hand-scheduled division of the binomial inversion loop, pda_device.cuh)
    against numpy's IEEE double division over the domain the loop can reach — and well beyond it."""
    rng = np.random.default_rng(14)
    n = 2_000_000
    # the recurrence itself: px' = ((n - X + 1) p px) / (X q)
    nn = rng.integers(1, 3000, n).astype(np.float64)
    X = np.minimum(rng.integers(1, 120, n), nn).astype(np.float64)
    p = np.concatenate([rng.uniform(1e-6, 0.5, n // 2), rng.uniform(0.0, 1.0, n - n // 2).astype(np.float32) / 2])
    px = np.exp(rng.uniform(-90.0, 0.0, n))
    num = (nn - X + 1.0) * p * px
    den = X * (1.0 - p)
    keep = num > 0
    num, den = num[keep], den[keep]
    # wide exponents, random mantissas
    m = 500_000
    num = np.concatenate([num, np.ldexp(rng.uniform(1.0, 2.0, m), rng.integers(-200, 60, m))])
    den = np.concatenate([den, np.ldexp(rng.uniform(1.0, 2.0, m), rng.integers(-40, 40, m))])
    # where the loop can really go with p near 1e-16: px falls by ~1e-16 per trip until the restart, the
    # numerator reaches ~1e-290 (2^-963) while the divisor X q stays within [0.5, 90]
    m = 500_000
    num = np.concatenate([num, np.ldexp(rng.uniform(1.0, 2.0, m), rng.integers(-963, -190, m))])
    den = np.concatenate([den, rng.uniform(0.5, 90.0, m)])
    Xs = rng.integers(1, 88, m).astype(np.float64)
    ps = np.concatenate([rng.uniform(1e-17, 1e-15, m // 2), rng.uniform(1e-12, 1e-6, m - m // 2)])
    num = np.concatenate([num, (rng.integers(1, 3000, m) + 1.0) * ps * np.ldexp(rng.uniform(1.0, 2.0, m), rng.integers(-910, 0, m))])
    den = np.concatenate([den, Xs * (1.0 - ps)])
    # mantissas next to a power of two and all-ones (the hard cases of reciprocal refinement)
    edge = np.array([1.0, 2.0 - 2.0 ** -52, 1.0 + 2.0 ** -52, 1.5, 2.0 - 2.0 ** -51, 1.0 + 2.0 ** -51, 4.0 / 3.0])
    en, ed = np.meshgrid(edge, edge)
    for s in (1.0, 3.0, 1e-30, 7.0, 2.0 ** -960):
        num = np.concatenate([num, en.ravel() * s]); den = np.concatenate([den, ed.ravel()])
    assert num.min() > 2.0 ** -1000 and np.isfinite(num).all() and (den > 0).all()
    k = len(num)
    got = mc1.probe(14, k, d0=np.stack([num, den], 1).reshape(-1), out="od")
    assert np.array_equal(got.view(np.uint64), (num / den).view(np.uint64))


def test_batch_matches_single_evaluations(gsm, oracle):
    w = synth.workload(4, n_bursts=700, R=8, P=12)
    b = w.bursts()
    with gsm.Mc(0, 2, keep_mce=True) as mc:
        mc.set_reSampleTimes(8); mc.set_seed(5); mc.set_hist(NB, LEFT_PAD)
        mc.init_bursts(b)
        mc.set_epoch(1, 40)
        mc.run_batch(b.n_bursts, 1, 2, w.args)
        mc.sync(1)
        chi = mc.chi2v(1)
        assert chi.shape == (12,)
        mcE = mc.hmcE(1, 12 * b.n_bursts * 8).reshape(12, -1)
        ohist = oracle.observed_hist(b, NB)
        for p in range(12):
            o = oracle.eval(b, 2, w.args[p], 8, 5, 40 + p)
            assert np.array_equal(bits(mcE[p]), bits(o["mcE"]))
            oh = oracle.hist(o["mcE"], NB)
            assert np.array_equal(mc.hist(1, p), oh)
            assert np.float32(chi[p]) == np.float32(oracle.chi2(ohist, oh, NB, 8, LEFT_PAD, 2))
        # resident re-run: epoch advanced by P
        mc.run_resident(b.n_bursts, 1); mc.sync(1)
        o = oracle.eval(b, 2, w.args[3], 8, 5, 52 + 3)
        assert np.array_equal(mc.hist(1, 3), oracle.hist(o["mcE"], NB))


@pytest.mark.parametrize("env", [
    {"GSM_CHUNK": "5", "GSM_SPAN": "3"}, {"GSM_CHUNK": "6", "GSM_SPAN": "32"}, {"GSM_CHUNK": "8", "GSM_SPAN": "2"},
    {"GSM_CHUNK": "8", "GSM_TAIL": "1"}, {"GSM_CHUNK": "7", "GSM_TAIL": "1", "GSM_OCC": "1"},
    {"GSM_CHUNK": "7", "GSM_TAIL": "0", "GSM_OCC": "0"}, {"GSM_OCC": "1"}, {"GSM_OCC": "0", "GSM_TAIL": "1"},
])
def test_results_do_not_depend_on_the_launch_shape(gsm, oracle, monkeypatch, env):
    """Chunk size, chunks per ticket (the span a warp draws for a population), the guided tail of small
    chunks and the register budget of the kernel build are launch-shape choices (pda_eval.cuh,
    shape_launch): every sample, histogram and chi-square must be the same bits whatever they are —
    for one vector (with background: both Poisson loops) and for a batch whose spans cross vector ends."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    w5 = synth.workload(5, n_bursts=1900, R=7)
    b5 = w5.bursts()
    with gsm.Mc(0, 1, keep_mce=True) as mc:
        mc.set_reSampleTimes(7); mc.set_seed(11); mc.set_hist(NB, LEFT_PAD)
        mc.init_bursts(b5)
        mc.set_epoch(0, 3)
        chi, h = mc.evaluate(4, w5.args[0], sid=0)
        o = oracle.eval(b5, 4, w5.args[0], 7, 11, 3)
        assert np.array_equal(bits(mc.hmcE(0, b5.n_bursts * 7)), bits(o["mcE"]))
        oh = oracle.hist(o["mcE"], NB)
        assert np.array_equal(h, oh)
        assert np.float32(chi) == np.float32(oracle.chi2(oracle.observed_hist(b5, NB), oh, NB, 7, LEFT_PAD, 4))
    w4 = synth.workload(4, n_bursts=411, R=5, P=9)
    b4 = w4.bursts()
    with gsm.Mc(0, 1, keep_mce=True) as mc:
        mc.set_reSampleTimes(5); mc.set_seed(12); mc.set_hist(NB, LEFT_PAD)
        mc.init_bursts(b4)
        mc.set_epoch(0, 70)
        mc.run_batch(b4.n_bursts, 0, 2, w4.args)
        mc.sync(0)
        mcE = mc.hmcE(0, 9 * b4.n_bursts * 5).reshape(9, -1)
        for p in range(9):
            o = oracle.eval(b4, 2, w4.args[p], 5, 12, 70 + p)
            assert np.array_equal(bits(mcE[p]), bits(o["mcE"]))
            assert np.array_equal(mc.hist(0, p), oracle.hist(o["mcE"], NB))


def test_streams_are_independent_and_async(gsm, oracle):
    w2 = synth.workload(2, n_bursts=1500, R=16)
    w3 = synth.workload(3, n_bursts=1500, R=16)
    b = w3.bursts()
    with gsm.Mc(0, 3) as mc:
        assert mc.streamNum == 3 and len(set(mc.getStreams())) == 3
        assert len(mc.gpuuuid) == 32
        mc.set_reSampleTimes(16); mc.set_seed(9); mc.set_hist(NB, LEFT_PAD)
        mc.init_bursts(b)
        for sid, (n, a) in enumerate([(2, w2.args[0]), (3, w3.args[0]), (2, w2.args[0])]):
            old = mc.set_nstates(n, sid)
            assert old == 0
            assert mc.set_params(n, sid, a)
            N = mc.setBurstBd(0, -1, sid)
            assert N == b.n_bursts
            mc.init_randstate(N, sid); mc.set_params_buff(old, N, sid)
            mc.set_epoch(sid, 100 + sid)
            mc.run_kernel(N, sid)
        for sid in range(3):
            while not mc.streamQuery(sid):
                pass
        for sid, (n, a) in enumerate([(2, w2.args[0]), (3, w3.args[0]), (2, w2.args[0])]):
            o = oracle.eval(b, n, a, 16, 9, 100 + sid)
            assert np.array_equal(mc.hist(sid, 0), oracle.hist(o["mcE"], NB))
        assert mc.set_nstates(3, 0) == 2           # previous n is returned
        assert not mc.streamQuery(7)               # out of range -> false, as mc::streamQuery


def test_setBurstBd_semantics(gsm):
    b = synth.workload(1, n_bursts=100).bursts()
    with gsm.Mc(0, 1) as mc:
        mc.init_bursts(b)
        assert mc.setBurstBd(0, -1, 0) == 100      # cstop < 0 -> sz_burst
        assert mc.setBurstBd(0, 1000, 0) == 100    # cstop >= sz_burst -> sz_burst
        assert mc.setBurstBd(10, 60, 0) == 50
        assert mc.setBurstBd(70, 60, 0) == 60      # cstart >= cstop -> start 0
        assert mc.setBurstBd(-1, 60, 0) == -1      # cstart == -1 -> N = -1


def test_edge_datasets(gsm, oracle):
    # empty dataset
    e = synth.make_bursts(0, (0.5,), 1)
    with gsm.Mc(0, 1) as mc:
        mc.set_hist(NB, LEFT_PAD)
        mc.init_bursts(e)
        chi2, hist = mc.evaluate(2, synth.pack_args([0.3, 0.7], [10, 10], [1, 1]))
        assert hist.sum() == 0
    # a burst starting at tick 0, photon-less bursts, zero durations, one burst only
    b = synth.make_bursts(40, (0.2, 0.8), 3, first_start=0)
    b.ph_count[[0, 5, 39]] = 0
    b.T[[2, 7]] = 0.0
    args = synth.pack_args([0.2, 0.8], [300, 700], [2, 2])
    for kw in ({}, dict(no_index=True), dict(force_raw64=True)):
        with gsm.Mc(0, 1, keep_mce=True, keep_counts=True, **kw) as mc:
            mc.set_reSampleTimes(5); mc.set_seed(1); mc.set_hist(NB, LEFT_PAD)
            mc.init_bursts(b)
            chi2, hist = mc.evaluate(2, args, epoch=2)
            o = oracle.eval(b, 2, args, 5, 1, 2)
            assert np.array_equal(bits(mc.hmcE(0, 200)), bits(o["mcE"]))
            assert np.array_equal(mc.steps(0, 200), o["n_steps"])
            assert np.array_equal(hist, oracle.hist(o["mcE"], NB))


def test_unpackable_ticks_fall_back_to_raw64(gsm, oracle):
    b = synth.make_bursts(50, (0.4,), 5)
    b.stop = b.stop + (1 << 31)                       # a 2^31-tick burst does not fit 30 bits
    last = b.istart + b.ph_count - 1
    b.burst_dd[last] = b.stop - 1
    b.burst_ad[last] = 0
    args = synth.pack_args([0.4, 0.6], [200, 200], [2, 2])
    with gsm.Mc(0, 1, keep_mce=True) as mc:
        mc.set_reSampleTimes(4); mc.set_seed(2); mc.set_hist(NB, LEFT_PAD)
        mc.init_bursts(b)
        assert mc.layout() == 0
        mc.evaluate(2, args, epoch=1)
        o = oracle.eval(b, 2, args, 4, 2, 1)
        assert np.array_equal(bits(mc.hmcE(0, 200)), bits(o["mcE"]))


def test_errors(gsm):
    b = synth.workload(1, n_bursts=50).bursts()
    with gsm.Mc(0, 1) as mc:
        with pytest.raises(gsm.GsmError):
            mc.run_kernel(10, 0)                        # before init_data
        mc.init_bursts(b)
        assert mc.set_params(2, 0, [0.3, 0.7, 1.0, 1.0, 2.0]) is False      # wrong count -> false
        assert mc.set_params(1, 0, [0.3, 2.0]) is False
        with pytest.raises(gsm.GsmError):
            mc.set_nstates(2, 5)                        # bad stream id
        with pytest.raises(gsm.GsmError):
            mc.hmcE(0, 10)                              # not kept
        bad = synth.workload(5, n_bursts=20).bursts()
        bad.bg_dd_rate = np.float32(5e6)                # lambda >= 1000: outside the Poisson contract
        with pytest.raises(gsm.GsmError):
            mc.init_bursts(bad)
        bad2 = synth.workload(1, n_bursts=20).bursts()
        bad2.istart = bad2.istart + 10 ** 9
        with pytest.raises(gsm.GsmError):
            mc.init_bursts(bad2)
    with pytest.raises(gsm.GsmError):
        gsm.Mc(99, 1)


def test_step_cap_flags_nonpositive_rates(gsm):
    b = synth.workload(1, n_bursts=64).bursts()
    with gsm.Mc(0, 1) as mc:
        mc.set_reSampleTimes(1); mc.set_hist(NB, LEFT_PAD)
        mc.init_bursts(b)
        mc.evaluate(2, synth.pack_args([0.3, 0.7], [-5.0, -5.0], [1, 1]))   # dwell times < 0: never ends
        assert mc.flags(0) & 1


@pytest.mark.parametrize("cfg", [2, 3, 5])
def test_full_size_properties(gsm, oracle, cfg):
    """BASELINE.json sizes (config 5 at 1e6 bursts: the generator, not the GPU, is the limit):
    histogram mass, determinism, epoch sensitivity, additivity over burst ranges, a bit-exact
    check of the first 2000 bursts sample by sample, and the whole histogram + chi-square of the
    full-size evaluation bit for bit against the oracle."""
    w = synth.workload(cfg, n_bursts=1_000_000 if cfg == 5 else None)
    b = w.bursts()
    B, R, n = b.n_bursts, w.R, w.n_states
    with gsm.Mc(0, 2) as mc:
        mc.set_reSampleTimes(R); mc.set_seed(77); mc.set_hist(NB, LEFT_PAD)
        mc.init_bursts(b)
        assert mc.layout() == 2
        c1, h1 = mc.evaluate(n, w.args[0], sid=0, epoch=5)
        c2, h2 = mc.evaluate(n, w.args[0], sid=1, epoch=5)
        c3, h3 = mc.evaluate(n, w.args[0], sid=0, epoch=6)
        assert int(h1.sum()) == B * R
        assert np.array_equal(h1, h2) and c1 == c2               # idempotent, stream-independent
        assert not np.array_equal(h1, h3)                         # a new epoch draws anew
        l1 = np.abs(h1.astype(np.int64) - h3.astype(np.int64)).sum() / (B * R)
        assert l1 < 0.02 and abs(c3 / c1 - 1) < 0.25               # ... from the same distribution
        _, ha = mc.evaluate(n, w.args[0], sid=0, N=2000, epoch=5)
        o = oracle.eval(b, n, w.args[0], R, 77, 5, 0, 2000)
        assert np.array_equal(ha, oracle.hist(o["mcE"], NB))
    # the WHOLE histogram and chi-square at full size, bit for bit, against the oracle on all host cores
    co, ho = oracle.eval_chi2(b, n, w.args[0], R, 77, 5, oracle.observed_hist(b, NB), NB, LEFT_PAD,
                              nthreads=len(os.sched_getaffinity(0)))
    assert np.array_equal(h1, ho), int((h1 != ho).sum())
    assert np.float32(c1).view(np.uint32) == np.float32(co).view(np.uint32)
    # additivity: two contexts holding the two halves with global burst ids sum to the whole
    half = B // 2
    hs = np.zeros(NB + 2, np.uint64)
    for (b0, b1) in ((0, half), (half, B)):
        with gsm.Mc(0, 1) as mc:
            mc.set_reSampleTimes(R); mc.set_seed(77); mc.set_hist(NB, LEFT_PAD)
            mc.init_bursts(b.slice(b0, b1), g_offset=b0)
            hs += mc.evaluate(n, w.args[0], epoch=5)[1]
    assert np.array_equal(hs, h1)


def test_coalesced_launch_equals_per_stream_launches(gsm, oracle):
    """gsm_run_coalesced: the vectors of several streams in one launch on a carrier slot, every
    vector with its own stream's epoch: each member then reads the bits its own launch would have
    produced; its epoch advanced by one; the next run of the member on its own slot is its own again."""
    w = synth.workload(4, n_bursts=900, R=8, P=5)
    b = w.bursts()
    ohist = oracle.observed_hist(b, NB)
    with gsm.Mc(0, 8) as mc:
        mc.set_reSampleTimes(8); mc.set_seed(5); mc.set_hist(NB, LEFT_PAD)
        mc.init_bursts(b)
        members = [1, 4, 2, 6, 0]
        for k, sid in enumerate(members):
            mc.set_epoch(sid, 100 * (k + 1))
        mc.run_coalesced(b.n_bursts, 7, 2, w.args, members)
        for k, sid in enumerate(members):
            while not mc.streamQuery(sid):
                pass
            c, h = oracle.eval_chi2(b, 2, w.args[k], 8, 5, 100 * (k + 1), ohist, NB, LEFT_PAD)
            assert np.array_equal(mc.hist(sid, 0), h)
            assert bits(np.float32(mc.chi2(sid))) == bits(np.float32(c))
            assert mc.chi2v(sid).shape == (1,)
            with pytest.raises(gsm.GsmError):
                mc.hist(sid, 1)                      # a member owns one vector of the carrier
            # the %globaltimer stamps a member reports are its carrier's launch, not its own stale ones
            t = mc.timing_ns(sid)
            assert t[0] > 0 and np.array_equal(t, mc.timing_ns(7))
        # the members' epochs moved on by one: the next run of a member is its own launch again
        c2, h2 = mc.evaluate(2, w.args[0], sid=4)
        co, ho = oracle.eval_chi2(b, 2, w.args[0], 8, 5, 201, ohist, NB, LEFT_PAD)
        assert np.array_equal(h2, ho) and bits(np.float32(c2)) == bits(np.float32(co))
        # refusals: the carrier among the members, a member twice
        with pytest.raises(gsm.GsmError):
            mc.run_coalesced(b.n_bursts, 7, 2, w.args[:2], [7, 1])
        with pytest.raises(gsm.GsmError):
            mc.run_coalesced(b.n_bursts, 7, 2, w.args[:2], [1, 1])
    with gsm.Mc(0, 3, keep_mce=True) as mc:
        mc.set_reSampleTimes(8); mc.set_hist(NB, LEFT_PAD)
        mc.init_bursts(b)
        with pytest.raises(gsm.GsmError):
            mc.run_coalesced(b.n_bursts, 2, 2, w.args[:2], [0, 1])   # per-sample outputs are per stream
