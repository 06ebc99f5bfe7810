"""Shared helpers of the test-suite."""
import glob
import os

import numpy as np

from gsmfretda_b200 import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
NB, LEFT_PAD = 200, 2


def golden_cases():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "*.npz")))


def load_golden(name):
    d = np.load(os.path.join(GOLDEN, name + ".npz"))
    b = synth.Bursts(start=d["start"], stop=d["stop"], istart=d["istart"], ph_count=d["ph_count"],
                     T=d["T"], burst_ad=d["burst_ad"], burst_dd=d["burst_dd"], SgDivSr=d["SgDivSr"],
                     clk_p=np.float32(d["clk_p"]), bg_ad_rate=np.float32(d["bg_ad_rate"]),
                     bg_dd_rate=np.float32(d["bg_dd_rate"]))
    return b, d


def bits(x):
    return np.ascontiguousarray(x, np.float32).view(np.uint32)


def r_squared(observed, expected):
    """Coefficient of determination of an observed pmf against the expected one
    (the role of dlib's r_squared in /root/reference/tests/gen_rand_tests.cpp)."""
    o = np.asarray(observed, np.float64)
    e = np.asarray(expected, np.float64)
    c = np.corrcoef(o, e)[0, 1]
    return float(c * c)


def ks_2samp_p(a, b):
    from scipy import stats
    return float(stats.ks_2samp(a, b).pvalue)
