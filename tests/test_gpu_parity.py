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
