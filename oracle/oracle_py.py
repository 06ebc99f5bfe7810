"""ctypes bindings of the CHECKER libraries — TEST INFRASTRUCTURE ONLY.

    libpda_oracle.so      the C restatement (oracle/pda_oracle.c)
    _ref/libref_cpu.so    the reference's own sources compiled for the host (oracle/ref_shim.cpp)
    _ref/libref_gpu.so    the reference's `class mc` unmodified for sm_100a (oracle/ref_gpu_shim.cpp)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
import this module.  Nothing under gsmfretda_b200/ does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
MAXS = 8

_f32p = C.POINTER(C.c_float)
_i64p = C.POINTER(C.c_int64)
_i32p = C.POINTER(C.c_int32)
_u32p = C.POINTER(C.c_uint32)
_u64p = C.POINTER(C.c_uint64)


class BurstsT(C.Structure):
    _fields_ = [("n_bursts", C.c_int64), ("start", _i64p), ("stop", _i64p), ("istart", _i64p),
                ("ph_count", _i32p), ("T", _f32p), ("burst_ad", _i64p), ("burst_dd", _i64p),
                ("clk_p", C.c_float), ("bg_ad_rate", C.c_float), ("bg_dd_rate", C.c_float)]


class ParamsT(C.Structure):
    _fields_ = [("n", C.c_int), ("E", C.c_float * MAXS), ("v", C.c_float * MAXS),
                ("K", C.c_float * (MAXS * MAXS)), ("P_i2j", C.c_float * (MAXS * MAXS)),
                ("p", C.c_float * MAXS)]


class SamplesT(C.Structure):
    _fields_ = [("mcE", _f32p), ("sum_ai", _i64p), ("F", _f32p), ("n_steps", _i32p)]


def build(ref: bool = True) -> None:
    """make the checker libraries (the reference build only where /root/reference exists)."""
    subprocess.run(["make", "-s", "-C", HERE, "all"], check=True)
    if ref and os.path.isdir("/root/reference/src"):
        subprocess.run(["make", "-s", "-C", HERE, "ref"], check=True)
        subprocess.run(["make", "-s", "-C", HERE, "refgpu"], check=True)


def _ptr(a, t):
    return a.ctypes.data_as(t)


class Oracle:
    def __init__(self):
        path = os.path.join(HERE, "libpda_oracle.so")
        if not os.path.exists(path):
            build(ref=False)
        L = self.lib = C.CDLL(path)
        L.pda_oracle_params.argtypes = [C.c_int, _f32p, C.c_int, C.POINTER(ParamsT)]
        L.pda_oracle_params.restype = C.c_int
        L.pda_oracle_eval.argtypes = [C.POINTER(BurstsT), C.POINTER(ParamsT), C.c_int, C.c_uint32,
                                      C.c_uint32, C.c_int64, C.c_int64, C.c_int64,
                                      C.POINTER(SamplesT), C.c_int]
        L.pda_oracle_eval.restype = None
        L.pda_oracle_hist.argtypes = [_f32p, C.c_int64, C.c_int, _u64p]
        L.pda_oracle_hist.restype = None
        L.pda_oracle_chi2.argtypes = [_f32p, _u64p, C.c_int, C.c_int, C.c_int, C.c_int]
        L.pda_oracle_chi2.restype = C.c_float
        L.pda_oracle_eval_chi2.argtypes = [C.POINTER(BurstsT), C.POINTER(ParamsT), C.c_int,
                                           C.c_uint32, C.c_uint32, C.c_int64, C.c_int64, C.c_int64,
                                           _f32p, C.c_int, C.c_int, _u64p, C.c_int]
        L.pda_oracle_eval_chi2.restype = C.c_float
        L.pda_oracle_philox.argtypes = [_u32p, _u32p, _u32p]
        L.pda_oracle_logf.argtypes = [C.c_float]; L.pda_oracle_logf.restype = C.c_float
        L.pda_oracle_log.argtypes = [C.c_double]; L.pda_oracle_log.restype = C.c_double
        L.pda_oracle_exp.argtypes = [C.c_double]; L.pda_oracle_exp.restype = C.c_double
        L.pda_oracle_norm_icdf.argtypes = [C.c_uint32]; L.pda_oracle_norm_icdf.restype = C.c_float
        L.pda_oracle_root6f.argtypes = [C.c_float]; L.pda_oracle_root6f.restype = C.c_float
        L.pda_oracle_u01.argtypes = [C.c_uint32]; L.pda_oracle_u01.restype = C.c_float
        L.pda_oracle_draw_dis_idx.argtypes = [C.c_int, _f32p, C.c_uint32]
        L.pda_oracle_draw_dis_idx.restype = C.c_int
        L.pda_oracle_draw_tau.argtypes = [C.c_float, C.c_uint32]; L.pda_oracle_draw_tau.restype = C.c_float
        L.pda_oracle_draw_e.argtypes = [C.c_float, C.c_float, C.c_float, C.c_uint32]
        L.pda_oracle_draw_e.restype = C.c_float
        L.pda_oracle_draw_j.argtypes = [_f32p, C.c_int, C.c_int, C.c_uint32]
        L.pda_oracle_draw_j.restype = C.c_int
        L.pda_oracle_poisson.argtypes = [C.c_uint32, C.c_uint32, C.c_double]
        L.pda_oracle_poisson.restype = C.c_uint32
        L.pda_oracle_rk_seed.argtypes = [C.c_uint64, _u32p]
        L.pda_oracle_rk_double.argtypes = [_u32p]; L.pda_oracle_rk_double.restype = C.c_double
        L.pda_oracle_binomial.argtypes = [_u32p, C.c_int, C.c_double]
        L.pda_oracle_binomial.restype = C.c_int64
        L.pda_oracle_multinomial.argtypes = [_u32p, C.c_int, C.c_int, _f32p, _i32p]
        L.pda_oracle_num_threads.restype = C.c_int

    # ---- structures ------------------------------------------------------------------
    @staticmethod
    def bursts_t(b) -> BurstsT:
        """b: gsmfretda_b200.synth.Bursts (arrays are kept alive by the caller)."""
        return BurstsT(b.n_bursts, _ptr(b.start, _i64p), _ptr(b.stop, _i64p), _ptr(b.istart, _i64p),
                       _ptr(b.ph_count, _i32p), _ptr(b.T, _f32p), _ptr(b.burst_ad, _i64p),
                       _ptr(b.burst_dd, _i64p), float(b.clk_p), float(b.bg_ad_rate),
                       float(b.bg_dd_rate))

    def params(self, n: int, args) -> ParamsT | None:
        a = np.ascontiguousarray(args, dtype=np.float32)
        out = ParamsT()
        ok = self.lib.pda_oracle_params(n, _ptr(a, _f32p), a.shape[0], C.byref(out))
        return out if ok else None

    # ---- evaluation ------------------------------------------------------------------
    def eval(self, b, n, args, R, seed, epoch, b0=0, b1=None, g_offset=0, nthreads=0):
        """-> dict(mcE f32, sum_ai i64, F f32, n_steps i32), each [(b1-b0)*R]."""
        b1 = b.n_bursts if b1 is None else b1
        tot = (b1 - b0) * R
        mcE = np.zeros(tot, np.float32); sa = np.zeros(tot, np.int64)
        F = np.zeros(tot, np.float32); ns = np.zeros(tot, np.int32)
        prm = self.params(n, args)
        assert prm is not None, "bad parameter vector"
        bt = self.bursts_t(b)
        st = SamplesT(_ptr(mcE, _f32p), _ptr(sa, _i64p), _ptr(F, _f32p), _ptr(ns, _i32p))
        self.lib.pda_oracle_eval(C.byref(bt), C.byref(prm), R, seed, epoch, b0, b1, g_offset,
                                 C.byref(st), nthreads)
        return dict(mcE=mcE, sum_ai=sa, F=F, n_steps=ns)

    def eval_sobol(self, b, n, args, R, evals=1, N=None, nthreads=0) -> np.ndarray:
        """The restatement on the reference's own generators (Sobol64 + xor128, persistent over
        `evals` evaluations) -> mcE[evals, N*R]: what the reference's class mc computes on the GPU."""
        N = b.n_bursts if N is None else N
        prm = self.params(n, args)
        assert prm is not None, "bad parameter vector"
        bt = self.bursts_t(b)
        dv, sc = sobol_tables()
        out = np.zeros((evals, N * R), np.float32)
        u64p = C.POINTER(C.c_uint64)
        self.lib.pda_oracle_eval_sobol.restype = None
        self.lib.pda_oracle_eval_sobol.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int64, u64p, u64p,
                                                   C.c_int, _f32p, C.c_int]
        self.lib.pda_oracle_eval_sobol(C.byref(bt), C.byref(prm), R, N, dv.ctypes.data_as(u64p),
                                       sc.ctypes.data_as(u64p), evals, out.ctypes.data_as(_f32p), nthreads)
        return out

    def hist(self, x, nb) -> np.ndarray:
        x = np.ascontiguousarray(x, np.float32)
        h = np.zeros(nb + 2, np.uint64)
        self.lib.pda_oracle_hist(_ptr(x, _f32p), x.shape[0], nb, _ptr(h, _u64p))
        return h

    def chi2(self, ohist, mhist, nb, R, left_pad, n_states) -> float:
        o = np.ascontiguousarray(ohist, np.float32)
        m = np.ascontiguousarray(mhist, np.uint64)
        return float(self.lib.pda_oracle_chi2(_ptr(o, _f32p), _ptr(m, _u64p), nb, R, left_pad, n_states))

    def eval_chi2(self, b, n, args, R, seed, epoch, ohist, nb, left_pad, b0=0, b1=None,
                  g_offset=0, nthreads=0):
        """-> (chi2, hist[nb+2]) without keeping per-sample outputs (cpu baseline leg)."""
        b1 = b.n_bursts if b1 is None else b1
        prm = self.params(n, args)
        assert prm is not None
        bt = self.bursts_t(b)
        o = np.ascontiguousarray(ohist, np.float32)
        h = np.zeros(nb + 2, np.uint64)
        c = self.lib.pda_oracle_eval_chi2(C.byref(bt), C.byref(prm), R, seed, epoch, b0, b1,
                                          g_offset, _ptr(o, _f32p), nb, left_pad, _ptr(h, _u64p),
                                          nthreads)
        return float(c), h

    def observed_hist(self, b, nb) -> np.ndarray:
        """streamWorker's ohist: counts of SgDivSr over nb bins, as float (streamWorker.cpp:62-70)."""
        return self.hist(b.SgDivSr, nb)[:nb].astype(np.float32)

    def num_threads(self) -> int:
        return int(self.lib.pda_oracle_num_threads())


class Ref:
    """The reference's own code (host build).  Exists only where /root/reference was present
    at build time; tests that need it skip otherwise."""

    def __init__(self):
        path = os.path.join(HERE, "_ref", "libref_cpu.so")
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        L = self.lib = C.CDLL(path)
        L.ref_params.argtypes = [C.c_int, _f32p, _f32p, _f32p, _f32p, _f32p, _f32p]
        L.ref_params.restype = C.c_int
        L.ref_genMatK.argtypes = [C.c_int, _f32p, C.c_int, _f32p]
        L.ref_genMatK.restype = C.c_int
        L.ref_mc_kernel.argtypes = [_i64p, _i64p, _i64p, _i64p, _i64p, _i32p, _f32p, C.c_float,
                                    C.c_float, C.c_float, _f32p, _f32p, _f32p, _f32p, _f32p,
                                    C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_uint32, C.c_int64,
                                    _f32p, C.c_int]
        L.ref_mc_kernel.restype = None
        for nm in ("ref_drawDisIdx",):
            getattr(L, nm).argtypes = [C.c_int, _f32p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
            getattr(L, nm).restype = C.c_int
        L.ref_drawTau.argtypes = [C.c_float, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
        L.ref_drawTau.restype = C.c_float
        L.ref_drawE.argtypes = [C.c_float, C.c_float, C.c_float, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
        L.ref_drawE.restype = C.c_float
        L.ref_drawJ.argtypes = [_f32p, C.c_int, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
        L.ref_drawJ.restype = C.c_int
        L.ref_rk_seed.argtypes = [C.c_uint64, _u32p]
        L.ref_rk_binomial.argtypes = [_u32p, C.c_int, C.c_double]
        L.ref_rk_binomial.restype = C.c_long
        L.ref_rk_double.argtypes = [_u32p]; L.ref_rk_double.restype = C.c_double
        L.ref_int_multinomial.argtypes = [_u32p, C.c_int, C.c_int, _f32p, _i32p]
        _u64p_ = C.POINTER(C.c_uint64)
        L.ref_mc_kernel_sobol.argtypes = [_i64p, _i64p, _i64p, _i64p, _i64p, _i32p, _f32p, C.c_float,
                                          C.c_float, C.c_float, _f32p, _f32p, _f32p, _f32p, _f32p,
                                          C.c_int, C.c_int, C.c_int, _u64p_, _u64p_, C.c_int, _f32p, C.c_int]
        L.ref_mc_kernel_sobol.restype = None

    def params(self, n, args):
        a = np.ascontiguousarray(args, np.float32).copy()
        E = np.zeros(n, np.float32); v = np.zeros(n, np.float32)
        K = np.zeros(n * n, np.float32); Pij = np.zeros(n * n, np.float32); p = np.zeros(n, np.float32)
        ok = self.lib.ref_params(n, _ptr(a, _f32p), _ptr(E, _f32p), _ptr(v, _f32p), _ptr(K, _f32p),
                                 _ptr(Pij, _f32p), _ptr(p, _f32p))
        return ok, E, v, K, Pij, p

    def mc_kernel(self, b, n, args, R, seed, epoch, N=None, g_offset=0, nthreads=0) -> np.ndarray:
        assert R & (R - 1) == 0, "the reference needs a power-of-two reSampleTimes (mc.cu:101)"
        N = b.n_bursts if N is None else N
        ok, E, v, K, Pij, p = self.params(n, args)
        assert ok
        out = np.zeros(N * R, np.float32)
        self.lib.ref_mc_kernel(_ptr(b.start, _i64p), _ptr(b.stop, _i64p), _ptr(b.burst_ad, _i64p),
                               _ptr(b.burst_dd, _i64p), _ptr(b.istart, _i64p), _ptr(b.ph_count, _i32p),
                               _ptr(b.T, _f32p), float(b.clk_p), float(b.bg_ad_rate),
                               float(b.bg_dd_rate), _ptr(E, _f32p), _ptr(v, _f32p), _ptr(K, _f32p),
                               _ptr(Pij, _f32p), _ptr(p, _f32p), N, n, R, seed, epoch, g_offset,
                               _ptr(out, _f32p), nthreads)
        return out


    def mc_kernel_sobol(self, b, n, args, R, evals=1, N=None, nthreads=0) -> np.ndarray:
        """The reference's setup_kernel + mc_kernel with cuRAND's own scrambled Sobol64 sequences
        (tables from libcurand's host API), `evals` evaluations on persistent RNG states as
        gpuWorker does -> mcE[evals, N*R]: what the reference's class mc computes on the GPU, up to
        libdevice-vs-det_math ulps."""
        assert R & (R - 1) == 0
        N = b.n_bursts if N is None else N
        ok, E, v, K, Pij, p = self.params(n, args)
        assert ok
        dv, sc = sobol_tables()
        out = np.zeros((evals, N * R), np.float32)
        u64p = C.POINTER(C.c_uint64)
        self.lib.ref_mc_kernel_sobol(_ptr(b.start, _i64p), _ptr(b.stop, _i64p), _ptr(b.burst_ad, _i64p),
                                     _ptr(b.burst_dd, _i64p), _ptr(b.istart, _i64p), _ptr(b.ph_count, _i32p),
                                     _ptr(b.T, _f32p), float(b.clk_p), float(b.bg_ad_rate),
                                     float(b.bg_dd_rate), _ptr(E, _f32p), _ptr(v, _f32p), _ptr(K, _f32p),
                                     _ptr(Pij, _f32p), _ptr(p, _f32p), N, n, R,
                                     dv.ctypes.data_as(u64p), sc.ctypes.data_as(u64p), evals,
                                     out.ctypes.data_as(_f32p), nthreads)
        return out


_SOBOL = None


def sobol_tables():
    """(direction vectors [20000, 64], scramble constants [20000]) of
    CURAND_SCRAMBLED_DIRECTION_VECTORS_64_JOEKUO6 from libcurand's host API — the tables
    mc::mc fetches (/root/reference/src/mc.cu:312-314).  Needs no GPU."""
    global _SOBOL
    if _SOBOL is None:
        import glob
        cands = (glob.glob("/usr/local/cuda/lib64/libcurand.so*") +
                 glob.glob(os.path.join(os.path.dirname(np.__file__), "..", "nvidia", "curand", "lib", "libcurand.so*")))
        if not cands:
            raise FileNotFoundError("libcurand.so")
        lib = C.CDLL(cands[0])
        p = C.POINTER(C.c_uint64)()
        q = C.POINTER(C.c_uint64)()
        assert lib.curandGetDirectionVectors64(C.byref(p), 104) == 0   # CURAND_SCRAMBLED_DIRECTION_VECTORS_64_JOEKUO6
        assert lib.curandGetScrambleConstants64(C.byref(q)) == 0
        _SOBOL = (np.ctypeslib.as_array(p, shape=(20000, 64)).copy(), np.ctypeslib.as_array(q, shape=(20000,)).copy())
    return _SOBOL


def have_ref() -> bool:
    return os.path.exists(os.path.join(HERE, "_ref", "libref_cpu.so"))


def have_refgpu() -> bool:
    return os.path.exists(os.path.join(HERE, "_ref", "libref_gpu.so"))


class RefGpu:
    """The reference's own `class mc` on the GPU (cuRAND Sobol64 + libdevice).  Its destructor
    calls cudaDeviceReset(): use it from a process of its own (oracle/refgpu_run.py) and never
    next to another CUDA user."""

    def __init__(self, devid=0, streams=1):
        path = os.path.join(HERE, "_ref", "libref_gpu.so")
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        L = self.lib = C.CDLL(path)
        L.refgpu_create.argtypes = [C.c_int, C.c_int]; L.refgpu_create.restype = C.c_void_p
        L.refgpu_destroy.argtypes = [C.c_void_p]
        L.refgpu_set_resample.argtypes = [C.c_void_p, C.c_int]
        L.refgpu_resample.argtypes = [C.c_void_p]; L.refgpu_resample.restype = C.c_int
        L.refgpu_init_data.argtypes = [C.c_void_p, C.c_int64, _i64p, _i64p, _i64p, _i32p, C.c_int64,
                                       _i64p, _i64p, _f32p, _f32p, C.c_float, C.c_float, C.c_float]
        L.refgpu_launch.argtypes = [C.c_void_p, C.c_int, _f32p, C.c_int]; L.refgpu_launch.restype = C.c_int
        L.refgpu_query.argtypes = [C.c_void_p, C.c_int]; L.refgpu_query.restype = C.c_int
        L.refgpu_hmcE.argtypes = [C.c_void_p, C.c_int]; L.refgpu_hmcE.restype = _f32p
        L.refgpu_time.argtypes = [C.c_void_p, C.c_int, _f32p, C.c_int, C.c_int]; L.refgpu_time.restype = C.c_double
        self.h = L.refgpu_create(devid, streams)
        self._keep = None

    def set_reSampleTimes(self, R):
        self.lib.refgpu_set_resample(self.h, R)
        return int(self.lib.refgpu_resample(self.h))

    def init_bursts(self, b):
        self._keep = b
        self.lib.refgpu_init_data(self.h, b.n_bursts, _ptr(b.istart, _i64p), _ptr(b.start, _i64p),
                                  _ptr(b.stop, _i64p), _ptr(b.ph_count, _i32p), b.burst_ad.shape[0],
                                  _ptr(b.burst_ad, _i64p), _ptr(b.burst_dd, _i64p), _ptr(b.T, _f32p),
                                  _ptr(b.SgDivSr, _f32p), float(b.clk_p), float(b.bg_ad_rate),
                                  float(b.bg_dd_rate))

    def evaluate(self, n, args, sid=0) -> np.ndarray:
        a = np.ascontiguousarray(args, np.float32)
        N = self.lib.refgpu_launch(self.h, n, _ptr(a, _f32p), sid)
        while not self.lib.refgpu_query(self.h, sid):
            pass
        R = int(self.lib.refgpu_resample(self.h))
        return np.ctypeslib.as_array(self.lib.refgpu_hmcE(self.h, sid), shape=(N * R,)).copy()

    def host_half(self, N, n, K, nb=200, left_pad=2, sid=0):
        """seconds for K host halves of an evaluation (vector copy + boost::histogram + chi-square,
        streamWorker.cpp:187-206) on the last hmcE; -> (seconds, chi2)"""
        self.lib.refgpu_host_half.restype = C.c_double
        self.lib.refgpu_host_half.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                              C.POINTER(C.c_float)]
        c = C.c_float(0)
        s = self.lib.refgpu_host_half(self.h, sid, N, nb, left_pad, n, K, C.byref(c))
        return float(s), float(c.value)

    def time(self, n, args, K, sid=0) -> float:
        a = np.ascontiguousarray(args, np.float32)
        return float(self.lib.refgpu_time(self.h, n, _ptr(a, _f32p), sid, K))
