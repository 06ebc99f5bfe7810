"""Golden vectors (tests/golden/*.npz, produced by tests/golden/make_golden.py from the
reference's own code): the oracle on CPU, and the CUDA path on the GPU."""
import numpy as np
import pytest

from helpers import NB, LEFT_PAD, bits, golden_cases, load_golden

CASES = golden_cases()


def test_fixtures_present():
    assert len(CASES) >= 9


@pytest.mark.parametrize("name", CASES)
def test_oracle_reproduces_reference_outputs(oracle, name):
    b, d = load_golden(name)
    n = int(d["n"])
    prm = oracle.params(n, d["args"])
    assert np.array_equal(bits(d["ref_K"]), bits(np.array(prm.K[:n * n])))
    assert np.array_equal(bits(d["ref_P_i2j"]), bits(np.array(prm.P_i2j[:n * n])))
    assert np.array_equal(bits(d["ref_p"]), bits(np.array(prm.p[:n])))
    o = oracle.eval(b, n, d["args"], int(d["R"]), int(d["seed"]), int(d["epoch"]))
    assert np.array_equal(bits(o["mcE"]), bits(d["ref_mcE"]))


@pytest.mark.gpu
@pytest.mark.parametrize("layout", ["indexed", "packed", "raw64"])
@pytest.mark.parametrize("name", CASES)
def test_cuda_reproduces_reference_outputs(gsm, oracle, name, layout):
    b, d = load_golden(name)
    n, R = int(d["n"]), int(d["R"])
    kw = dict(indexed={}, packed=dict(no_index=True), raw64=dict(force_raw64=True))[layout]
    with gsm.Mc(0, 1, keep_mce=True, **kw) as mc:
        mc.set_reSampleTimes(R)
        mc.set_seed(int(d["seed"]))
        mc.set_hist(NB, LEFT_PAD)
        mc.init_bursts(b)
        chi2, hist = mc.evaluate(n, d["args"], sid=0, epoch=int(d["epoch"]))
        mcE = mc.hmcE(0, b.n_bursts * R)
        assert mc.layout() == dict(indexed=2, packed=1, raw64=0)[layout]
    assert np.array_equal(bits(mcE), bits(d["ref_mcE"]))
    oh = oracle.hist(d["ref_mcE"], NB)
    assert np.array_equal(hist, oh)
    c = oracle.chi2(oracle.observed_hist(b, NB), oh, NB, R, LEFT_PAD, n)
    assert np.float32(chi2).view(np.uint32) == np.float32(c).view(np.uint32) or (np.isnan(c) and np.isnan(chi2))
