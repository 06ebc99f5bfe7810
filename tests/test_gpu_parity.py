"""GPU parity: the CUDA path, called through the C ABI, against the oracle on the same
seeded inputs.  Bar: bit-exact per-sample efficiencies, integer acceptor counts, dwell
steps, histogram bins and chi-square (replay mode: identical draws)."""
import numpy as np
import pytest

from gsmfretda_b200 import synth

pytestmark = pytest.mark.gpu

NB, LEFT_PAD = 200, 2


def run_gpu(gsm, b, n, args, R, seed, epoch, N=None, **kw):
    with gsm.Mc(0, 2, keep_mce=True, keep_counts=True, **kw) as mc:
        mc.set_reSampleTimes(R)
        mc.set_seed(seed)
        mc.set_hist(NB, LEFT_PAD)
        mc.init_bursts(b)
        chi2, hist = mc.evaluate(n, args, sid=0, N=N, epoch=epoch)
        Ns = b.n_bursts if N is None else N
        tot = Ns * R
        return dict(chi2=chi2, hist=hist, mcE=mc.hmcE(0, tot), sum_ai=mc.counts(0, tot),
                    n_steps=mc.steps(0, tot), packed=mc.is_packed(), layout=mc.layout(), flags=mc.flags(0),
                    ohist=mc.observed_hist())


def check(gsm, oracle, b, n, args, R, seed=11, epoch=3, N=None, **kw):
    g = run_gpu(gsm, b, n, args, R, seed, epoch, N=N, **kw)
    Ns = b.n_bursts if N is None else N
    o = oracle.eval(b, n, args, R, seed, epoch, 0, Ns)
    assert g["flags"] == 0
    assert np.array_equal(g["n_steps"], o["n_steps"])
    assert np.array_equal(g["sum_ai"], o["sum_ai"])
    assert np.array_equal(g["mcE"].view(np.uint32), o["mcE"].view(np.uint32))
    oh = oracle.hist(o["mcE"], NB)
    assert np.array_equal(g["hist"], oh)
    assert int(g["hist"].sum()) == Ns * R
    ohist = oracle.observed_hist(b, NB)
    assert np.array_equal(g["ohist"], ohist)
    c = oracle.chi2(ohist, oh, NB, R, LEFT_PAD, n)
    assert np.float32(g["chi2"]).view(np.uint32) == np.float32(c).view(np.uint32) or \
        (np.isnan(c) and np.isnan(g["chi2"]))
    return g


@pytest.mark.parametrize("cfg,nb,R", [(1, 3000, 4), (1, 777, 64), (2, 2000, 100), (3, 3000, 16),
                                      (5, 3000, 4), (2, 1, 1), (3, 257, 7)])
def test_configs_bit_exact(gsm, oracle, cfg, nb, R):
    w = synth.workload(cfg, n_bursts=nb, R=R)
    g = check(gsm, oracle, w.bursts(), w.n_states, w.args[0], R)
    assert g["packed"] and g["layout"] == 2


@pytest.mark.parametrize("cfg,nb,R", [(1, 1000, 4), (2, 1500, 100), (3, 2000, 16), (5, 2000, 4)])
def test_packed_linear_scan_bit_exact(gsm, oracle, cfg, nb, R):
    w = synth.workload(cfg, n_bursts=nb, R=R)
    g = check(gsm, oracle, w.bursts(), w.n_states, w.args[0], R, no_index=True)
    assert g["layout"] == 1


@pytest.mark.parametrize("cfg", [2, 5])
def test_raw64_layout_bit_exact(gsm, oracle, cfg):
    w = synth.workload(cfg, n_bursts=1500, R=8)
    g = check(gsm, oracle, w.bursts(), w.n_states, w.args[0], 8, force_raw64=True)
    assert not g["packed"] and g["layout"] == 0


def test_partial_range(gsm, oracle):
    w = synth.workload(3, n_bursts=1000, R=16)
    check(gsm, oracle, w.bursts(), w.n_states, w.args[0], 16, N=333)


def many_state_case(n, bg, nb=1500, seed=9100):
    """n states, full K matrix (every off-diagonal rate its own), optionally with background."""
    rng = np.random.default_rng(seed + n)
    E = np.linspace(0.08, 0.92, n)
    k = np.exp(rng.uniform(np.log(150.0), np.log(4000.0), size=n * n - n))
    args = synth.pack_args(E, k, rng.uniform(1.0, 3.0, size=n))
    b = synth.make_bursts(nb, E, seed=seed + 7 * n, bg_ad_rate=1200.0 * bg, bg_dd_rate=2000.0 * bg)
    return b, args


@pytest.mark.parametrize("n,bg", [(5, 0), (5, 1), (6, 0), (6, 1), (7, 0), (8, 0), (8, 1)])
def test_many_states_bit_exact(gsm, oracle, n, bg):
    """n = 5, 6 are the top of the reference's own test matrix (/root/reference/tests/gen_rand_tests.cpp:79-82);
    n = 8 is the top of get_max_conversion_capacity's range (/root/reference/src/mc.cu:402-438)."""
    b, args = many_state_case(n, bg)
    for kw in ({}, dict(no_index=True)):
        g = check(gsm, oracle, b, n, args, 8, seed=31 + n, epoch=n, **kw)
        assert g["packed"]


def test_many_states_batch_bit_exact(gsm, oracle):
    """A batch of 6-state vectors in one launch equals the oracle vector by vector."""
    n = 6
    b, a0 = many_state_case(n, 1, nb=800)
    rng = np.random.default_rng(5)
    vecs = np.stack([a0 * np.concatenate([np.ones(n), rng.uniform(0.5, 2.0, n * n - n), np.ones(n)]).astype(np.float32)
                     for _ in range(5)])
    with gsm.Mc(0, 1) as mc:
        mc.set_reSampleTimes(4); mc.set_seed(99); mc.set_hist(NB, LEFT_PAD)
        mc.init_bursts(b)
        mc.set_epoch(0, 40)
        mc.run_batch(b.n_bursts, 0, n, vecs)
        mc.sync(0)
        chi = mc.chi2v(0)
        ohist = oracle.observed_hist(b, NB)
        for p in range(5):
            c, h = oracle.eval_chi2(b, n, vecs[p], 4, 99, 40 + p, ohist, NB, LEFT_PAD)
            assert np.array_equal(mc.hist(0, p), h)
            assert np.float32(chi[p]).view(np.uint32) == np.float32(c).view(np.uint32)
