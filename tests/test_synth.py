"""The synthetic burst generator: deterministic, and in the layout mc::init_data_gpu takes."""
import numpy as np

from gsmfretda_b200 import synth


def test_deterministic_and_well_formed():
    a = synth.workload(3, n_bursts=500).bursts()
    b = synth.workload(3, n_bursts=500).bursts()
    for f in ("start", "stop", "istart", "ph_count", "T", "burst_ad", "burst_dd", "SgDivSr"):
        assert np.array_equal(getattr(a, f), getattr(b, f))
    assert a.n_photons == int(a.ph_count.sum())
    assert np.all(np.diff(a.start) > 0) and np.all(a.stop > a.start)
    assert np.all(a.ph_count >= 21) and np.all(a.ph_count <= 300)
    assert np.array_equal(a.istart, np.concatenate([[0], np.cumsum(a.ph_count)[:-1]]))
    t = np.maximum(a.burst_ad, a.burst_dd)
    for i in (0, 17, 499):
        s = slice(int(a.istart[i]), int(a.istart[i] + a.ph_count[i]))
        real = t[s][t[s] > 0]
        assert np.all(np.diff(real) >= 0) and real.min() >= a.start[i] and real.max() <= a.stop[i]
    assert not np.any((a.burst_ad > 0) & (a.burst_dd > 0))
    assert np.all((a.SgDivSr >= 0) & (a.SgDivSr <= 1))
    assert a.algorithmic_bytes() == 36 * 500 + 16 * a.n_photons


def test_slice_rebases_istart():
    a = synth.workload(2, n_bursts=300).bursts()
    s = a.slice(100, 200)
    assert s.n_bursts == 100 and s.istart[0] == 0
    assert np.array_equal(s.burst_ad, a.burst_ad[a.istart[100]:a.istart[200]])


def test_arg_layout():
    w = synth.workload(3)
    assert w.args.shape == (1, 12)
    w4 = synth.workload(4)
    assert w4.args.shape == (1024, 6) and w4.P == 1024
    assert np.all((w4.args[:, :2] > 0.05) & (w4.args[:, :2] < 0.95))
