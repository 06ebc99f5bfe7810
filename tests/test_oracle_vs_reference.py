"""The C restatement (oracle/pda_oracle.c) against the reference's own sources built for the
host (oracle/_ref/libref_cpu.so) on the same draws: bit-exact.  Skipped where the reference
build is absent (it is made where /root/reference exists)."""
import ctypes as C

import numpy as np
import pytest

from gsmfretda_b200 import synth
from helpers import bits


@pytest.mark.parametrize("n", [2, 3, 4, 5, 6])
def test_params_match_eigenhelper(oracle, ref, n):
    rng = np.random.default_rng(n)
    for _ in range(20):
        # MatCheck::genRandMatk ranges: E in [0,1), k in [0,1000), v in [0,10)  (tests/ParamsTest.cpp:10-24)
        args = np.concatenate([rng.random(n), rng.random(n * n - n) * 1000, rng.random(n) * 10]).astype(np.float32)
        ok, E, v, K, Pij, p = ref.params(n, args)
        prm = oracle.params(n, args)
        assert ok and prm is not None
        assert np.array_equal(bits(E), bits(np.array(prm.E[:n])))
        assert np.array_equal(bits(v), bits(np.array(prm.v[:n])))
        assert np.array_equal(bits(K), bits(np.array(prm.K[:n * n])))
        assert np.array_equal(bits(Pij), bits(np.array(prm.P_i2j[:n * n])))
        assert np.array_equal(bits(p), bits(np.array(prm.p[:n])))


def test_genMatK_rejects_like_the_reference(oracle, ref):
    K = np.zeros(9, np.float32)
    k5 = np.ones(5, np.float32)
    assert ref.lib.ref_genMatK(3, k5.ctypes.data_as(C.POINTER(C.c_float)), 5, K.ctypes.data_as(C.POINTER(C.c_float))) == 0
    assert oracle.params(3, np.ones(11, np.float32)) is None


def test_draw_primitives_match(oracle, ref):
    rng = np.random.default_rng(5)
    L, R = oracle.lib, ref.lib
    f32p = C.POINTER(C.c_float)
    for trial in range(300):
        g, rep, seed, epoch = (int(x) for x in rng.integers(0, 2 ** 32, 4))
        ctr = (C.c_uint32 * 4)(g, rep, 0, 0); key = (C.c_uint32 * 2)(seed, epoch); w = (C.c_uint32 * 4)()
        L.pda_oracle_philox(ctr, key, w)
        n = int(rng.integers(2, 7))
        p = rng.random(n).astype(np.float32)
        p /= p.sum()
        # drawDisIdx consumes word x0 in phase 1 of the shim
        assert L.pda_oracle_draw_dis_idx(n, p.ctypes.data_as(f32p), w[0]) == \
            R.ref_drawDisIdx(n, p.ctypes.data_as(f32p), g, rep, seed, epoch)
        k = float(rng.random() * 5000 + 1e-3)
        a, b = L.pda_oracle_draw_tau(k, w[1]), R.ref_drawTau(k, g, rep, seed, epoch)
        assert np.float32(a).view(np.uint32) == np.float32(b).view(np.uint32)
        e, v = float(rng.uniform(0.02, 0.98)), float(rng.uniform(0, 6))
        a, b = L.pda_oracle_draw_e(e, 52.0, v, w[2]), R.ref_drawE(e, 52.0, v, g, rep, seed, epoch)
        assert np.float32(a).view(np.uint32) == np.float32(b).view(np.uint32)
        P = rng.random((n, n)).astype(np.float32)
        i = int(rng.integers(0, n))
        Pf = np.ascontiguousarray(P.reshape(-1))
        assert L.pda_oracle_draw_j(Pf.ctypes.data_as(f32p), n, i, w[0]) == \
            R.ref_drawJ(Pf.ctypes.data_as(f32p), n, i, g, rep, seed, epoch)


def test_xor128_and_binomial_match(oracle, ref):
    rng = np.random.default_rng(6)
    L, R = oracle.lib, ref.lib
    for trial in range(400):
        s64 = int(rng.integers(0, 2 ** 63))
        a = (C.c_uint32 * 4)(); b = (C.c_uint32 * 4)()
        L.pda_oracle_rk_seed(s64, a); R.ref_rk_seed(s64, b)
        assert list(a) == list(b)
        assert L.pda_oracle_rk_double(a) == R.ref_rk_double(b)
        # inversion (n*p <= 30) and BTPE (n*p > 30; never covered by the reference's tests)
        n = int(rng.choice([0, 1, 2, 3, 5, 8, 20, 70, 150, 400, 3000]))
        p = float(rng.choice([1e-3, 0.05, 0.3, 0.5, 0.62, 0.9, 0.999]))
        for _ in range(5):
            assert L.pda_oracle_binomial(a, n, p) == R.ref_rk_binomial(b, n, p)
        assert list(a) == list(b)


def test_multinomial_matches(oracle, ref):
    rng = np.random.default_rng(7)
    L, R = oracle.lib, ref.lib
    for trial in range(100):
        K = int(rng.integers(2, 7))
        p = rng.random(K).astype(np.float32)
        N = int(rng.integers(1, 500))
        a = (C.c_uint32 * 4)(); b = (C.c_uint32 * 4)()
        L.pda_oracle_rk_seed(trial + 1, a); R.ref_rk_seed(trial + 1, b)
        oa = np.zeros(K, np.int32); ob = np.zeros(K, np.int32)
        L.pda_oracle_multinomial(a, K, N, p.ctypes.data_as(C.POINTER(C.c_float)), oa.ctypes.data_as(C.POINTER(C.c_int32)))
        R.ref_int_multinomial(b, K, N, p.ctypes.data_as(C.POINTER(C.c_float)), ob.ctypes.data_as(C.POINTER(C.c_int32)))
        assert np.array_equal(oa, ob) and oa.sum() <= N


@pytest.mark.parametrize("cfg,nb,R", [(1, 1500, 4), (2, 1200, 16), (3, 1500, 8), (4, 800, 4), (5, 1500, 4)])
def test_mc_kernel_matches(oracle, ref, cfg, nb, R):
    w = synth.workload(cfg, n_bursts=nb, R=R, P=2 if cfg == 4 else None)
    b = w.bursts()
    for pi in range(w.P):
        o = oracle.eval(b, w.n_states, w.args[pi], R, 31 + cfg, pi)
        r = ref.mc_kernel(b, w.n_states, w.args[pi], R, 31 + cfg, pi)
        assert np.array_equal(bits(o["mcE"]), bits(r))


def test_mc_kernel_burst_offset_and_range(oracle, ref):
    w = synth.workload(3, n_bursts=600, R=4)
    b = w.bursts()
    o = oracle.eval(b, 3, w.args[0], 4, 9, 9, 0, 250, g_offset=1000)
    r = ref.mc_kernel(b, 3, w.args[0], 4, 9, 9, N=250, g_offset=1000)
    assert np.array_equal(bits(o["mcE"]), bits(r))
