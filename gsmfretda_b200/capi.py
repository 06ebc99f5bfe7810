"""ctypes binding of libgsmfret.so (include/gsmfret_c.h) and `Mc`, a Python mirror of the
reference's `class mc` (/root/reference/src/mc.hpp:78-179) on top of it.

The library is the product; there is no fallback.  If libgsmfret.so is missing or no sm_100
device is present, construction raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GSMFRET_LIB") or os.path.join(_HERE, "libgsmfret.so")

GSM_KEEP_MCE = 1
GSM_KEEP_COUNTS = 2
GSM_FORCE_RAW64 = 4
GSM_NO_INDEX = 8
GSM_NO_TIMING = 16
GSM_PEER_SHARED_DEVICE = 32
GSM_MODE_REPLAY = 0
GSM_MODE_NORMAL = 1
GSM_MODE_REFERENCE = 2
PEER_HANDLE_BYTES = 128
GSM_FLAG_STEP_CAP = 1
GSM_FLAG_PEER_TIMEOUT = 2

_f32p = C.POINTER(C.c_float)
_f64p = C.POINTER(C.c_double)
_i64p = C.POINTER(C.c_int64)
_i32p = C.POINTER(C.c_int32)
_u32p = C.POINTER(C.c_uint32)
_u64p = C.POINTER(C.c_uint64)
_ctx = C.c_void_p

# name -> (restype, argtypes): every symbol include/gsmfret_c.h declares
SIGNATURES = {
    "gsm_abi_version": (C.c_int, []),
    "gsm_last_error": (C.c_char_p, []),
    "gsm_device_count": (C.c_int, []),
    "gsm_create": (C.c_int, [C.c_int, C.c_int, C.c_uint, C.POINTER(_ctx)]),
    "gsm_destroy": (None, [_ctx]),
    "gsm_profiler": (C.c_int, [C.c_int]),
    "gsm_set_device": (C.c_int, [_ctx]),
    "gsm_stream_count": (C.c_int, [_ctx]),
    "gsm_stream": (C.c_void_p, [_ctx, C.c_int]),
    "gsm_device_id": (C.c_int, [_ctx]),
    "gsm_device_uuid": (C.c_int, [_ctx, C.c_char_p]),
    "gsm_set_resample_times": (C.c_int, [_ctx, C.c_int]),
    "gsm_resample_times": (C.c_int, [_ctx]),
    "gsm_set_seed": (C.c_int, [_ctx, C.c_uint32]),
    "gsm_set_mode": (C.c_int, [_ctx, C.c_int]),
    "gsm_mode": (C.c_int, [_ctx]),
    "gsm_set_hist": (C.c_int, [_ctx, C.c_int, C.c_int]),
    "gsm_init_data": (C.c_int, [_ctx, C.c_int64, _i64p, _i64p, _i64p, _i32p, C.c_int64, _i64p, _i64p,
                                _f32p, _f32p, C.c_float, C.c_float, C.c_float, C.c_int64]),
    "gsm_n_bursts": (C.c_int64, [_ctx]),
    "gsm_is_packed": (C.c_int, [_ctx]),
    "gsm_layout": (C.c_int, [_ctx]),
    "gsm_observed_hist": (C.c_int, [_ctx, _f32p, C.c_int]),
    "gsm_set_observed_hist": (C.c_int, [_ctx, _f32p, C.c_int]),
    # device pointers and a cudaStream_t, as integers
    "gsm_observed_hist_static": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p]),
    "gsm_get_max_conversion_capacity": (C.c_int, [_ctx, C.c_int]),
    "gsm_set_nstates": (C.c_int, [_ctx, C.c_int, C.c_int]),
    "gsm_set_params": (C.c_int, [_ctx, C.c_int, C.c_int, _f32p, C.c_int]),
    "gsm_set_burst_bd": (C.c_int, [_ctx, C.c_int, C.c_int, C.c_int]),
    "gsm_init_randstate": (C.c_int, [_ctx, C.c_int, C.c_int]),
    "gsm_set_params_buff": (C.c_int, [_ctx, C.c_int, C.c_int, C.c_int]),
    "gsm_set_epoch": (C.c_int, [_ctx, C.c_int, C.c_uint32]),
    "gsm_run": (C.c_int, [_ctx, C.c_int, C.c_int]),
    "gsm_run_batch": (C.c_int, [_ctx, C.c_int, C.c_int, C.c_int, C.c_int, _f32p]),
    "gsm_run_coalesced": (C.c_int, [_ctx, C.c_int, C.c_int, C.c_int, C.c_int, _f32p, _i32p]),
    "gsm_run_resident": (C.c_int, [_ctx, C.c_int, C.c_int]),
    "gsm_query": (C.c_int, [_ctx, C.c_int]),
    "gsm_sync": (C.c_int, [_ctx, C.c_int]),
    "gsm_result_chi2": (C.c_float, [_ctx, C.c_int]),
    "gsm_result_chi2v": (_f32p, [_ctx, C.c_int, C.POINTER(C.c_int)]),
    "gsm_result_hist": (_u64p, [_ctx, C.c_int, C.c_int]),
    "gsm_result_mcE": (_f32p, [_ctx, C.c_int]),
    "gsm_result_counts": (_i64p, [_ctx, C.c_int]),
    "gsm_result_steps": (_i32p, [_ctx, C.c_int]),
    "gsm_result_flags": (C.c_uint64, [_ctx, C.c_int]),
    "gsm_result_timing": (C.c_int, [_ctx, C.c_int, _u64p]),
    "gsm_last_elapsed_ms": (C.c_float, [_ctx, C.c_int]),
    "gsm_launch_count": (C.c_uint64, [_ctx]),
    "gsm_comm_unique_id": (C.c_int, [C.c_void_p]),
    "gsm_comm_init": (C.c_int, [_ctx, C.c_void_p, C.c_int, C.c_int]),
    "gsm_comm_destroy": (C.c_int, [_ctx]),
    "gsm_peer_export": (C.c_int, [_ctx, C.c_int, C.c_int, C.c_void_p]),
    "gsm_peer_connect": (C.c_int, [_ctx, C.c_void_p, C.c_int, C.c_int]),
    "gsm_peer_disconnect": (C.c_int, [_ctx]),
    "gsm_peer_set_timeout": (C.c_int, [_ctx, C.c_uint32]),
    "gsm_probe": (C.c_int, [_ctx, C.c_int, C.c_int64, C.c_int, C.c_int, _u32p, _u32p, _f32p, _f32p,
                            _f64p, _i32p, _u32p, _f32p, _f64p, _i64p, _i32p]),
    "gsm_issue_peak": (C.c_int, [_ctx, C.c_int, C.POINTER(C.c_double)]),
    "gsm_launch_shape": (C.c_int, [C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int64)]),
}

_lib = None


def load() -> C.CDLL:
    """dlopen libgsmfret.so and type every entry point.  Raises if the library is absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(make -C gsmfretda_b200/csrc). There is no CPU fallback.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


class GsmError(RuntimeError):
    pass


def _p(a, t):
    return None if a is None else a.ctypes.data_as(t)


class Mc:
    """Python mirror of `class mc`; method names follow the reference.

    Extra (fused path): `hist(sid, p)`, `chi2(sid)`; `hmcE(sid)` needs keep_mce=True."""

    def __init__(self, devid: int = 0, streamNum: int = 16, keep_mce: bool = False,
                 keep_counts: bool = False, force_raw64: bool = False, no_index: bool = False,
                 peer_shared_device: bool = False):
        self.lib = load()
        flags = (GSM_KEEP_MCE if keep_mce else 0) | (GSM_KEEP_COUNTS if keep_counts else 0) | \
                (GSM_FORCE_RAW64 if force_raw64 else 0) | (GSM_NO_INDEX if no_index else 0) | \
                (GSM_PEER_SHARED_DEVICE if peer_shared_device else 0)
        h = _ctx()
        self._h = None
        self._ck(self.lib.gsm_create(devid, streamNum, flags, C.byref(h)))
        self._h = h
        self.devid = devid
        self.streamNum = self.lib.gsm_stream_count(h)
        self.s_n = [0] * self.streamNum
        self._keep = []
        self.sz_burst = 0

    # -- plumbing ---------------------------------------------------------------------
    def _ck(self, rc: int) -> int:
        if rc < 0:
            raise GsmError(f"gsmfret error {rc}: {self.lib.gsm_last_error().decode()}")
        return rc

    def close(self):
        if self._h is not None:
            self.lib.gsm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- the reference's members ------------------------------------------------------
    def set_gpuid(self):
        self._ck(self.lib.gsm_set_device(self._h))

    def getStreams(self):
        return [self.lib.gsm_stream(self._h, i) for i in range(self.streamNum)]

    @property
    def gpuuuid(self) -> str:
        buf = C.create_string_buffer(33)
        self._ck(self.lib.gsm_device_uuid(self._h, buf))
        return buf.value.decode()

    @property
    def reSampleTimes(self) -> int:
        return self.lib.gsm_resample_times(self._h)

    def set_reSampleTimes(self, n: int, exact: bool = True):
        """exact=False applies the reference's rounding down to a power of two (mc.cu:316)."""
        if not exact:
            n = 1 << (int(n).bit_length() - 1)
        self._ck(self.lib.gsm_set_resample_times(self._h, int(n)))

    def set_seed(self, seed: int):
        self._ck(self.lib.gsm_set_seed(self._h, seed & 0xFFFFFFFF))

    def set_mode(self, mode: int):
        """GSM_MODE_REPLAY (bit-exact draw contract) or GSM_MODE_NORMAL (FP32 fast arithmetic)."""
        self._ck(self.lib.gsm_set_mode(self._h, int(mode)))

    def mode(self) -> int:
        return int(self.lib.gsm_mode(self._h))

    def set_hist(self, n_bins: int = 200, left_pad: int = 2):
        self._ck(self.lib.gsm_set_hist(self._h, n_bins, left_pad))
        self.n_bins = n_bins

    def init_data_gpu(self, istart, start, stop, phCount, sz_tag, burst_ad, burst_dd, T, SgDivSr,
                      clk_p, bg_ad_rate, bg_dd_rate, g_offset: int = 0):
        a = lambda x, dt: np.ascontiguousarray(x, dtype=dt)
        istart, start, stop = a(istart, np.int64), a(start, np.int64), a(stop, np.int64)
        phCount = a(phCount, np.int32)
        burst_ad, burst_dd = a(burst_ad, np.int64), a(burst_dd, np.int64)
        T, SgDivSr = a(T, np.float32), a(SgDivSr, np.float32)
        B = start.shape[0]
        self._ck(self.lib.gsm_init_data(self._h, B, _p(istart, _i64p), _p(start, _i64p), _p(stop, _i64p),
                                        _p(phCount, _i32p), int(sz_tag), _p(burst_ad, _i64p),
                                        _p(burst_dd, _i64p), _p(T, _f32p), _p(SgDivSr, _f32p),
                                        float(clk_p), float(bg_ad_rate), float(bg_dd_rate), int(g_offset)))
        self.sz_burst = B
        if not hasattr(self, "n_bins"):
            self.n_bins = 200

    def init_bursts(self, b, g_offset: int = 0):
        """init_data_gpu from a synth.Bursts."""
        self.init_data_gpu(b.istart, b.start, b.stop, b.ph_count, b.n_photons, b.burst_ad, b.burst_dd,
                           b.T, b.SgDivSr, b.clk_p, b.bg_ad_rate, b.bg_dd_rate, g_offset)

    def get_max_conversion_capacity(self, max_stateNum: int = 8) -> bool:
        return bool(self.lib.gsm_get_max_conversion_capacity(self._h, max_stateNum))

    def set_nstates(self, n: int, sid: int) -> int:
        r = self._ck(self.lib.gsm_set_nstates(self._h, n, sid))
        self.s_n[sid] = n
        return r

    def set_params(self, n: int, sid: int, args) -> bool:
        a = np.ascontiguousarray(args, dtype=np.float32)
        return bool(self._ck(self.lib.gsm_set_params(self._h, n, sid, _p(a, _f32p), a.shape[0])))

    def setBurstBd(self, cstart: int, cstop: int, sid: int) -> int:
        r = self.lib.gsm_set_burst_bd(self._h, cstart, cstop, sid)
        if r < -1:
            self._ck(r)
        return r

    def init_randstate(self, N: int, sid: int):
        self._ck(self.lib.gsm_init_randstate(self._h, N, sid))

    def set_params_buff(self, oldS_n: int, N_sid: int, sid: int):
        self._ck(self.lib.gsm_set_params_buff(self._h, oldS_n, N_sid, sid))

    def run_kernel(self, N: int, sid: int):
        self._ck(self.lib.gsm_run(self._h, N, sid))

    def streamQuery(self, sid: int) -> bool:
        return self._ck(self.lib.gsm_query(self._h, sid)) == 1

    # -- extensions -------------------------------------------------------------------
    def set_epoch(self, sid: int, epoch: int):
        self._ck(self.lib.gsm_set_epoch(self._h, sid, epoch & 0xFFFFFFFF))

    def run_batch(self, N: int, sid: int, n: int, args2d):
        a = np.ascontiguousarray(args2d, dtype=np.float32)
        assert a.ndim == 2 and a.shape[1] == n * n + n
        self._ck(self.lib.gsm_run_batch(self._h, N, sid, n, a.shape[0], _p(a, _f32p)))

    def run_coalesced(self, N: int, carrier_sid: int, n: int, args2d, member_sids):
        """The single vectors of several streams in one launch on `carrier_sid`; afterwards the
        members' streamQuery / chi2 / hist read their vector of the carrier."""
        a = np.ascontiguousarray(args2d, dtype=np.float32)
        m = np.ascontiguousarray(member_sids, dtype=np.int32)
        assert a.ndim == 2 and a.shape[1] == n * n + n and m.shape[0] == a.shape[0]
        self._ck(self.lib.gsm_run_coalesced(self._h, N, carrier_sid, n, a.shape[0], _p(a, _f32p), _p(m, _i32p)))

    def run_resident(self, N: int, sid: int):
        self._ck(self.lib.gsm_run_resident(self._h, N, sid))

    def sync(self, sid: int):
        self._ck(self.lib.gsm_sync(self._h, sid))

    def chi2(self, sid: int) -> float:
        return float(self.lib.gsm_result_chi2(self._h, sid))

    def chi2v(self, sid: int) -> np.ndarray:
        P = C.c_int(0)
        ptr = self.lib.gsm_result_chi2v(self._h, sid, C.byref(P))
        if not ptr:
            raise GsmError("no results on this stream")
        return np.ctypeslib.as_array(ptr, shape=(P.value,)).copy()

    def hist(self, sid: int, p: int = 0) -> np.ndarray:
        ptr = self.lib.gsm_result_hist(self._h, sid, p)
        if not ptr:
            raise GsmError("no results on this stream")
        return np.ctypeslib.as_array(ptr, shape=(self.n_bins + 2,)).copy()

    def hmcE(self, sid: int, count: int) -> np.ndarray:
        ptr = self.lib.gsm_result_mcE(self._h, sid)
        if not ptr:
            raise GsmError("per-sample efficiencies were not kept (keep_mce=False)")
        return np.ctypeslib.as_array(ptr, shape=(count,)).copy()

    def counts(self, sid: int, count: int) -> np.ndarray:
        ptr = self.lib.gsm_result_counts(self._h, sid)
        if not ptr:
            raise GsmError("per-sample counts were not kept (keep_counts=False)")
        return np.ctypeslib.as_array(ptr, shape=(count,)).copy()

    def steps(self, sid: int, count: int) -> np.ndarray:
        ptr = self.lib.gsm_result_steps(self._h, sid)
        if not ptr:
            raise GsmError("per-sample steps were not kept (keep_counts=False)")
        return np.ctypeslib.as_array(ptr, shape=(count,)).copy()

    def flags(self, sid: int) -> int:
        return int(self.lib.gsm_result_flags(self._h, sid))

    def timing_ns(self, sid: int) -> np.ndarray:
        """%globaltimer stamps of the last run's last block: start, posted, all peers seen, done."""
        out = (C.c_uint64 * 4)()
        self._ck(self.lib.gsm_result_timing(self._h, sid, out))
        return np.array(list(out), dtype=np.uint64)

    def stream_ptr(self, sid: int) -> int:
        """The slot's cudaStream_t as an integer (torch.cuda.ExternalStream takes it)."""
        return int(self.lib.gsm_stream(self._h, sid) or 0)

    def elapsed_ms(self, sid: int) -> float:
        return float(self.lib.gsm_last_elapsed_ms(self._h, sid))

    def issue_peak(self, kind: int = 0) -> float:
        """Measured thread-level multiply-adds per second (0: integer, 1: FP32) on this GPU."""
        v = C.c_double(0.0)
        self._ck(self.lib.gsm_issue_peak(self._h, kind, C.byref(v)))
        return float(v.value)

    def launch_count(self) -> int:
        return int(self.lib.gsm_launch_count(self._h))

    def is_packed(self) -> bool:
        return bool(self.lib.gsm_is_packed(self._h))

    def layout(self) -> int:
        return int(self.lib.gsm_layout(self._h))

    def observed_hist(self) -> np.ndarray:
        out = np.zeros(self.n_bins, np.float32)
        self._ck(self.lib.gsm_observed_hist(self._h, _p(out, _f32p), self.n_bins))
        return out

    def set_observed_hist(self, ohist):
        o = np.ascontiguousarray(ohist, np.float32)
        self._ck(self.lib.gsm_set_observed_hist(self._h, _p(o, _f32p), o.shape[0]))

    def comm_init(self, id128: bytes, rank: int, world: int):
        buf = C.create_string_buffer(bytes(id128), 128)
        self._ck(self.lib.gsm_comm_init(self._h, buf, rank, world))

    def comm_destroy(self):
        self._ck(self.lib.gsm_comm_destroy(self._h))

    def peer_export(self, world: int, max_P: int = 1) -> bytes:
        """Allocate this rank's symmetric exchange block; returns the 128-byte handle to gather."""
        buf = C.create_string_buffer(PEER_HANDLE_BYTES)
        self._ck(self.lib.gsm_peer_export(self._h, world, max_P, buf))
        return buf.raw

    def peer_connect(self, handles, rank: int, world: int):
        """handles: the world handles in rank order (bytes of world x 128, or a list of bytes)."""
        blob = handles if isinstance(handles, (bytes, bytearray)) else b"".join(handles)
        assert len(blob) == world * PEER_HANDLE_BYTES, (len(blob), world)
        buf = C.create_string_buffer(bytes(blob), len(blob))
        self._ck(self.lib.gsm_peer_connect(self._h, buf, rank, world))

    def peer_disconnect(self):
        self._ck(self.lib.gsm_peer_disconnect(self._h))

    def peer_set_timeout(self, ms: int):
        self._ck(self.lib.gsm_peer_set_timeout(self._h, ms))

    def evaluate(self, n: int, args, sid: int = 0, N: int | None = None, epoch: int | None = None):
        """The six calls gpuWorker makes per evaluation (gpuWorker.cpp:71-80), then wait.
        -> (chi2, hist[n_bins+2])"""
        self.set_nstates(n, sid)
        if not self.set_params(n, sid, args):
            raise GsmError("set_params rejected the vector")
        Ns = self.setBurstBd(0, -1 if N is None else N, sid)
        if epoch is not None:
            self.set_epoch(sid, epoch)
        self.run_kernel(Ns, sid)
        self.sync(sid)
        return self.chi2(sid), self.hist(sid, 0)

    def probe(self, what: int, n: int, ni: int = 0, nj: int = 0, w0=None, w1=None, f0=None, f1=None,
              d0=None, i0=None, out: str = "of", out_len: int | None = None):
        c = lambda x, dt: None if x is None else np.ascontiguousarray(x, dtype=dt)
        w0, w1 = c(w0, np.uint32), c(w1, np.uint32)
        f0, f1, d0, i0 = c(f0, np.float32), c(f1, np.float32), c(d0, np.float64), c(i0, np.int32)
        m = n if out_len is None else out_len
        outs = dict(ou=None, of=None, od=None, ol=None, oi=None)
        dt = dict(ou=np.uint32, of=np.float32, od=np.float64, ol=np.int64, oi=np.int32)[out]
        outs[out] = np.zeros(m, dt)
        self._ck(self.lib.gsm_probe(self._h, what, n, ni, nj, _p(w0, _u32p), _p(w1, _u32p), _p(f0, _f32p),
                                    _p(f1, _f32p), _p(d0, _f64p), _p(i0, _i32p), _p(outs["ou"], _u32p),
                                    _p(outs["of"], _f32p), _p(outs["od"], _f64p), _p(outs["ol"], _i64p),
                                    _p(outs["oi"], _i32p)))
        return outs[out]


def comm_unique_id() -> bytes:
    lib = load()
    buf = C.create_string_buffer(128)
    rc = lib.gsm_comm_unique_id(buf)
    if rc < 0:
        raise GsmError(f"gsmfret error {rc}: {lib.gsm_last_error().decode()}")
    return buf.raw
