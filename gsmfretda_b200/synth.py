"""Synthetic photon-burst generator (numpy only).

Produces the arrays `mc::init_data_gpu` takes (/root/reference/src/mc.cu:456-521), i.e. what
`loadhdf5` + `burst_data` (/root/reference/src/loadHdf5.cpp:17-128) would hand over, for the
five BASELINE.json configurations.  The recipe is SURVEY.md §8(d):

    phCount_b = 20 + min(280, Geometric(mean 40));  T_b = phCount_b / 30 kHz
    start_b strictly increasing with 1 ms gaps, stop_b = start_b + T_b/clk_p (80 MHz ticks)
    photon ticks sorted uniform in [start_b, stop_b]; a photon is acceptor-emission w.p.
    0.8*E_b (burst_ad = t, burst_dd = 0), donor-emission w.p. 0.8*(1-E_b) (burst_dd = t),
    else both 0 (acceptor-excitation photon); SgDivSr_b = n_ad / (n_ad + n_dd).

The real fixture tests/data/sample.hdf5 of the reference is a Git-LFS pointer and is
unavailable on this box, so config 1 is "fixture-sized" synthetic data.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

CLK_P = np.float32(1.25e-8)


@dataclass
class Bursts:
    """Burst arrays in the reference's layout (all host numpy, C-contiguous)."""

    start: np.ndarray      # int64 [B]
    stop: np.ndarray       # int64 [B]
    istart: np.ndarray     # int64 [B] first photon of burst b in burst_ad/burst_dd
    ph_count: np.ndarray   # int32 [B]
    T: np.ndarray          # float32 [B] burst duration, seconds
    burst_ad: np.ndarray   # int64 [Nph] timestamp if acceptor-emission photon else 0
    burst_dd: np.ndarray   # int64 [Nph] timestamp if donor-emission photon else 0
    SgDivSr: np.ndarray    # float32 [B] observed proximity ratio
    clk_p: np.float32 = CLK_P
    bg_ad_rate: np.float32 = np.float32(0.0)
    bg_dd_rate: np.float32 = np.float32(0.0)

    @property
    def n_bursts(self) -> int:
        return int(self.start.shape[0])

    @property
    def n_photons(self) -> int:
        return int(self.burst_ad.shape[0])

    def algorithmic_bytes(self) -> int:
        """Bytes of burst data one evaluation has to read in the reference's layout:
        36 B per burst + 16 B per photon."""
        return 36 * self.n_bursts + 16 * self.n_photons

    def slice(self, b0: int, b1: int) -> "Bursts":
        """Bursts [b0, b1) with istart rebased (burst-split sharding)."""
        p0 = int(self.istart[b0]) if b0 < self.n_bursts else self.n_photons
        p1 = int(self.istart[b1]) if b1 < self.n_bursts else self.n_photons
        return Bursts(
            start=self.start[b0:b1].copy(), stop=self.stop[b0:b1].copy(),
            istart=(self.istart[b0:b1] - p0), ph_count=self.ph_count[b0:b1].copy(),
            T=self.T[b0:b1].copy(), burst_ad=self.burst_ad[p0:p1].copy(),
            burst_dd=self.burst_dd[p0:p1].copy(), SgDivSr=self.SgDivSr[b0:b1].copy(),
            clk_p=self.clk_p, bg_ad_rate=self.bg_ad_rate, bg_dd_rate=self.bg_dd_rate)


def make_bursts(n_bursts: int, E_states, seed: int, bg_ad_rate: float = 0.0,
                bg_dd_rate: float = 0.0, state_weights=None, first_start: int = 1_000_000,
                max_extra: int = 280, mean_extra: float = 40.0) -> Bursts:
    """Deterministic synthetic bursts; each burst sits in one ground-truth state."""
    rng = np.random.default_rng(seed)
    B = int(n_bursts)
    E_states = np.asarray(E_states, dtype=np.float64)
    if state_weights is None:
        state_weights = np.full(E_states.shape[0], 1.0 / E_states.shape[0])
    ph = 20 + np.minimum(max_extra, rng.geometric(1.0 / mean_extra, size=B))
    ph = ph.astype(np.int32)
    T = (ph.astype(np.float64) / 3.0e4).astype(np.float32)
    dur = np.floor(T.astype(np.float64) / float(CLK_P)).astype(np.int64)
    gap = np.int64(round(1e-3 / float(CLK_P)))
    # start_b = first_start + sum_{j<b} (dur_j + gap)
    start = np.empty(B, dtype=np.int64)
    if B:
        start[0] = first_start
        np.cumsum(dur[:-1] + gap, out=start[1:])
        start[1:] += first_start
    stop = start + dur
    istart = np.zeros(B, dtype=np.int64)
    if B:
        np.cumsum(ph[:-1].astype(np.int64), out=istart[1:])
    nph = int(ph.sum())
    burst_of = np.repeat(np.arange(B, dtype=np.int64), ph)
    # sorted uniforms per burst by normalised exponential spacings (O(Nph), no sort)
    e = rng.exponential(size=nph)
    tail = rng.exponential(size=B)
    csum = np.cumsum(e)
    base = np.zeros(B)
    if B:
        base[1:] = csum[istart[1:] - 1]
    within = csum - base[burst_of]
    tot = (np.add.reduceat(e, istart) if B else np.zeros(0)) + tail
    frac = within / tot[burst_of]
    t = start[burst_of] + np.floor(frac * dur[burst_of]).astype(np.int64)
    Eb = E_states[rng.choice(E_states.shape[0], size=B, p=state_weights)]
    u = rng.random(nph)
    pa = 0.8 * Eb[burst_of]
    is_a = u < pa
    is_d = (~is_a) & (u < 0.8)
    burst_ad = np.where(is_a, t, 0).astype(np.int64)
    burst_dd = np.where(is_d, t, 0).astype(np.int64)
    n_a = np.add.reduceat(is_a.astype(np.int64), istart) if B else np.zeros(0, np.int64)
    n_d = np.add.reduceat(is_d.astype(np.int64), istart) if B else np.zeros(0, np.int64)
    with np.errstate(invalid="ignore", divide="ignore"):
        sg = np.where(n_a + n_d > 0, n_a / np.maximum(n_a + n_d, 1), 0.0).astype(np.float32)
    return Bursts(start=start, stop=stop, istart=istart, ph_count=ph, T=T,
                  burst_ad=burst_ad, burst_dd=burst_dd, SgDivSr=sg, clk_p=CLK_P,
                  bg_ad_rate=np.float32(bg_ad_rate), bg_dd_rate=np.float32(bg_dd_rate))


def concat(blocks) -> Bursts:
    """One dataset from several blocks of bursts, in order (istart rebased block by block).  The
    blocks keep their own time axes: nothing downstream needs start_b to increase across blocks."""
    blocks = list(blocks)
    if len(blocks) == 1:
        return blocks[0]
    off, ist = 0, []
    for b in blocks:
        ist.append(b.istart + off)
        off += b.n_photons
    cat = lambda name: np.concatenate([getattr(b, name) for b in blocks])
    b0 = blocks[0]
    return Bursts(start=cat("start"), stop=cat("stop"), istart=np.concatenate(ist), ph_count=cat("ph_count"),
                  T=cat("T"), burst_ad=cat("burst_ad"), burst_dd=cat("burst_dd"), SgDivSr=cat("SgDivSr"),
                  clk_p=b0.clk_p, bg_ad_rate=b0.bg_ad_rate, bg_dd_rate=b0.bg_dd_rate)


def pack_args(E, k_offdiag_rowmajor, v) -> np.ndarray:
    """args = [E_0..E_{n-1} | k (n*n-n, row-major without the diagonal) | v_0..v_{n-1}],
    the vector `mc::set_params` splits (/root/reference/src/mc.cu:792-799)."""
    return np.concatenate([np.asarray(E, np.float32), np.asarray(k_offdiag_rowmajor, np.float32),
                           np.asarray(v, np.float32)]).astype(np.float32)


@dataclass
class Workload:
    name: str
    n_bursts: int
    R: int
    n_states: int
    args: np.ndarray            # [P, n*n+n] float32
    E_states: tuple
    bg_ad_rate: float = 0.0
    bg_dd_rate: float = 0.0
    seed: int = 0
    notes: str = ""
    extra: dict = field(default_factory=dict)

    @property
    def P(self) -> int:
        return int(self.args.shape[0])

    def bursts(self, n_bursts: int | None = None, seed: int | None = None) -> Bursts:
        return make_bursts(self.n_bursts if n_bursts is None else n_bursts, self.E_states,
                           self.seed if seed is None else seed, self.bg_ad_rate, self.bg_dd_rate)


def workload(cfg: int, n_bursts: int | None = None, R: int | None = None,
             P: int | None = None) -> Workload:
    """The five BASELINE.json configurations."""
    if cfg == 1:
        w = Workload("cfg1_fixture_sized_2state_static", 7733, 4, 2,
                     pack_args([0.3, 0.7], [1e-3, 1e-3], [2, 2])[None, :], (0.3, 0.7), seed=1)
    elif cfg == 2:
        w = Workload("cfg2_2state_dynamic_1e5x100", 100_000, 100, 2,
                     pack_args([0.25, 0.75], [500, 500], [3, 3])[None, :], (0.25, 0.75), seed=2)
    elif cfg == 3:
        k = [1000, 2500, 1800, 5000, 3200, 1400]
        w = Workload("cfg3_3state_fullK_1e6x16", 1_000_000, 16, 3,
                     pack_args([0.2, 0.5, 0.8], k, [2, 2, 2])[None, :], (0.2, 0.5, 0.8), seed=3)
    elif cfg == 4:
        rng = np.random.default_rng(4)
        n_p = 1024 if P is None else P
        E = rng.uniform(0.05, 0.95, size=(n_p, 2))
        k = np.exp(rng.uniform(np.log(10.0), np.log(1e4), size=(n_p, 2)))
        v = rng.uniform(0.5, 5.0, size=(n_p, 2))
        args = np.concatenate([E, k, v], axis=1).astype(np.float32)
        w = Workload("cfg4_population1024_2state_1e5x16", 100_000, 16, 2, args, (0.3, 0.7), seed=4)
    elif cfg == 5:
        k = [200, 650, 1100, 1550, 2000, 380, 830, 1280, 1730, 560, 1010, 1460]
        w = Workload("cfg5_4state_bg_1e7x4", 10_000_000, 4, 4,
                     pack_args([0.15, 0.4, 0.65, 0.9], k, [2, 2, 2, 2])[None, :],
                     (0.15, 0.4, 0.65, 0.9), bg_ad_rate=1500.0, bg_dd_rate=2500.0, seed=5)
    else:
        raise ValueError(f"unknown config {cfg}")
    if n_bursts is not None:
        w.n_bursts = int(n_bursts)
    if R is not None:
        w.R = int(R)
    if P is not None and cfg != 4:
        w.args = np.repeat(w.args, P, axis=0)
    return w


def save_bin(b: Bursts, path: str) -> None:
    """Flat binary dump read by gsmfretda_b200/host/local_node.cpp (format in its header)."""
    import struct
    with open(path, "wb") as f:
        f.write(b"GSMB")
        f.write(struct.pack("<Iqq4f", 1, b.n_bursts, b.n_photons, float(b.clk_p), float(b.bg_ad_rate),
                            float(b.bg_dd_rate), 0.0))
        for a, dt in ((b.start, np.int64), (b.stop, np.int64), (b.istart, np.int64), (b.ph_count, np.int32),
                      (b.T, np.float32), (b.SgDivSr, np.float32), (b.burst_ad, np.int64), (b.burst_dd, np.int64)):
            np.ascontiguousarray(a, dtype=dt).tofile(f)
