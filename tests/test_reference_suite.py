"""The reference's own test matrix re-hosted:
  * tests/ParamsTest.cpp:26-71   algebraic pins of genMatK / genP_i2j / genMatP, n in 2..6
  * tests/gen_rand_tests.cpp + GenRandTest.cu  statistical pins of the draw primitives, same
    thresholds (R^2 >= 0.6 / 0.711, multinomial moments within 30 %), with numpy standing in
    for GSL and dlib.
CPU variants run on the oracle; GPU variants run the device code through gsm_probe."""
import ctypes as C

import numpy as np
import pytest

from helpers import r_squared

N_SET = [1000, 10000, 20000, 30000]          # gen_rand_tests.cpp:79
F_SET = [0.41, 0.54, 2.1, 3.3, 334.3, 4343, 72343]


@pytest.mark.parametrize("n", [2, 3, 4, 5, 6])
def test_params_algebra(oracle, n):
    rng = np.random.default_rng(100 + n)
    args = np.concatenate([rng.random(n), rng.random(n * n - n) * 1000, rng.random(n) * 10]).astype(np.float32)
    prm = oracle.params(n, args)
    K = np.array(prm.K[:n * n]).reshape(n, n).T      # column-major -> K[i, j]
    P = np.array(prm.P_i2j[:n * n]).reshape(n, n).T
    # Map(0,i) == matK[i*n]  (ParamsTest.cpp:57): column-major storage
    for i in range(n):
        assert K[0, i] == prm.K[i * n]
    # every column of P_i2j sums to 1 +- 1e-5 (ParamsTest.cpp:60-62)
    assert np.allclose(P.sum(axis=0), 1.0, atol=1e-5)
    assert np.all(np.diag(P) == 0)
    # off-diagonal ratios of P equal those of K (ParamsTest.cpp:63-66, n = 3 in the reference)
    for j in range(n):
        rows = [i for i in range(n) if i != j]
        for a in rows:
            for b in rows:
                if K[b, j] > 0:
                    assert abs(P[a, j] / P[b, j] - K[a, j] / K[b, j]) < 1e-4 * max(1.0, K[a, j] / K[b, j])
    # diagonal of K = minus the column's off-diagonal sum; p = row inflow, not normalised
    for j in range(n):
        assert abs(K[j, j] + (K[:, j].sum() - K[j, j])) < 1e-3 * abs(K[j, j])
    assert np.allclose(np.array(prm.p[:n]), K.sum(axis=1) - np.diag(K), rtol=1e-5)


def _words(rng, n):
    return rng.integers(0, 2 ** 32, n, dtype=np.uint64).astype(np.uint32)


def _cpu_probe(oracle):
    L = oracle.lib
    f32p = C.POINTER(C.c_float)

    class P:
        def categorical(self, p, w):
            p = np.ascontiguousarray(p, np.float32)
            return np.array([L.pda_oracle_draw_dis_idx(p.size, p.ctypes.data_as(f32p), int(x)) for x in w])

        def tau(self, k, w):
            return np.array([L.pda_oracle_draw_tau(float(k), int(x)) for x in w], np.float32)

        def binomial(self, n, p, seeds):
            out = []
            st = (C.c_uint32 * 4)()
            for s in seeds:
                L.pda_oracle_rk_seed(int(s), st)
                out.append(L.pda_oracle_binomial(st, n, float(p)))
            return np.array(out)

        def multinomial(self, p, N, seeds):
            p = np.ascontiguousarray(p, np.float32)
            out = np.zeros((len(seeds), p.size), np.int32)
            st = (C.c_uint32 * 4)()
            for i, s in enumerate(seeds):
                L.pda_oracle_rk_seed(int(s), st)
                L.pda_oracle_multinomial(st, p.size, N, p.ctypes.data_as(f32p), out[i].ctypes.data_as(C.POINTER(C.c_int32)))
            return out
    return P()


class _GpuProbe:
    def __init__(self, gsm):
        self.mc = gsm.Mc(0, 1)

    def categorical(self, p, w):
        return self.mc.probe(12, len(w), ni=len(p), w0=w, f0=p, out="oi")

    def tau(self, k, w):
        return self.mc.probe(10, len(w), w0=w, f0=np.full(len(w), k, np.float32), out="of")

    def binomial(self, n, p, seeds):
        s = np.asarray(seeds, np.uint64)
        w = np.stack([(s & 0xffffffff).astype(np.uint32), (s >> 32).astype(np.uint32)], 1).reshape(-1)
        return self.mc.probe(7, len(seeds), ni=1, w0=w, i0=np.full(len(seeds), n, np.int32),
                             d0=np.full(len(seeds), p, np.float64), out="ol")

    def multinomial(self, p, N, seeds):
        s = np.asarray(seeds, np.uint64)
        w = np.stack([(s & 0xffffffff).astype(np.uint32), (s >> 32).astype(np.uint32)], 1).reshape(-1)
        K = len(p)
        return self.mc.probe(11, len(seeds), ni=K, nj=N, w0=w, f0=p, out="oi", out_len=len(seeds) * K).reshape(-1, K)


def _stat_suite(pr, N, n, f):
    rng = np.random.default_rng(int(N + 31 * n + f * 7) % (2 ** 32))
    # drawDisIdx: R^2 >= 0.6 against the target pmf (GenRandTest.cu:218-221)
    p = rng.random(n).astype(np.float32)
    p /= p.sum()
    idx = pr.categorical(p, _words(rng, N))
    obs = np.bincount(idx, minlength=n) / N
    assert r_squared(obs, p) >= 0.6
    # drawTau(k = f): 100-bin histogram against the exponential law, R^2 >= 0.711 for N > 10000
    # (GenRandTest.cu:263-266)
    if N > 10000:
        t = pr.tau(f, _words(rng, N)).astype(np.float64)
        ref = rng.exponential(1.0 / f, N)
        hi = max(t.max(), ref.max())
        h1, _ = np.histogram(t, bins=100, range=(0, hi))
        h2, _ = np.histogram(ref, bins=100, range=(0, hi))
        assert r_squared(h1 / N, h2 / N) >= 0.711
    # drawJ_Si2Sj: column of P_i2j without the diagonal = categorical over n-1 (GenRandTest.cu:170-173)
    col = rng.random(n - 1).astype(np.float32)
    col /= col.sum()
    if n - 1 >= 2:
        j = pr.categorical(col, _words(rng, N))
        assert r_squared(np.bincount(j, minlength=n - 1) / N, col) >= 0.711
    # rk_binomial(n, p): histogram against a library binomial (gsl_ran_binomial there), R^2 >= 0.711
    # (GenRandTest.cu:347-350).  The reference uses n <= 6 (inversion only); BTPE is added below.
    pb = float(min(0.95, max(0.05, (f % 1.0) or 0.37)))
    seeds = rng.integers(1, 2 ** 62, N)
    x = pr.binomial(n, pb, seeds)
    ref = rng.binomial(n, pb, N)
    assert r_squared(np.bincount(x, minlength=n + 1) / N, np.bincount(ref, minlength=n + 1) / N) >= 0.711
    # csd_int_multinomial: mean / variance / covariance within 30 % of theory (GenRandTest.cu:404-418)
    if N >= 20000:
        pm = rng.random(n).astype(np.float32) + 0.2
        Ntr = 50
        m = pr.multinomial(pm, Ntr, seeds[:8000]).astype(np.float64)
        q = pm / pm.sum()
        assert np.all(m.sum(axis=1) <= Ntr)
        assert np.allclose(m.mean(axis=0), Ntr * q, rtol=0.3)
        assert np.allclose(m.var(axis=0), Ntr * q * (1 - q), rtol=0.3)
        cov = np.cov(m.T)
        for a in range(n):
            for b in range(a + 1, n):
                assert abs(cov[a, b] + Ntr * q[a] * q[b]) <= 0.3 * Ntr * q[a] * q[b] + 0.15


@pytest.mark.parametrize("N,n,f", [(1000, 3, 0.41), (20000, 4, 3.3), (30000, 6, 4343)])
def test_genrand_statistics_oracle(oracle, N, n, f):
    _stat_suite(_cpu_probe(oracle), N, n, f)


def test_btpe_statistics_oracle(oracle):
    rng = np.random.default_rng(8)
    pr = _cpu_probe(oracle)
    for n, p in ((200, 0.4), (1000, 0.5), (90, 0.62)):
        x = pr.binomial(n, p, rng.integers(1, 2 ** 62, 20000))
        ref = rng.binomial(n, p, 20000)
        assert r_squared(np.bincount(x, minlength=n + 1), np.bincount(ref, minlength=n + 1)) >= 0.95
        assert abs(x.mean() - n * p) < 0.1 and abs(x.var() / (n * p * (1 - p)) - 1) < 0.06


@pytest.mark.gpu
@pytest.mark.parametrize("N", N_SET)
@pytest.mark.parametrize("n", [3, 4, 5, 6])
@pytest.mark.parametrize("f", F_SET)
def test_genrand_statistics_gpu(gsm, N, n, f):
    pr = _GpuProbe(gsm)
    try:
        _stat_suite(pr, N, n, f)
    finally:
        pr.mc.close()
