"""Host-side planning of a burst-split evaluation across GPUs.

Every rank holds a contiguous range of the bursts, balanced by photon count, and keys its
draws by the GLOBAL burst index (`g_offset`), so the summed integer histogram — and with it
chi-square — is bit-identical whatever the number of ranks.  The reference has nothing of the
kind: its process-per-GPU launcher (/root/reference/src/call_main.cpp:195-227) can only hand
different parameter vectors to different GPUs through the server.
"""
from __future__ import annotations

import numpy as np


def split_bursts(ph_count, world: int):
    """-> list of (b0, b1) contiguous ranges, one per rank, balanced by sum(ph_count) + 36-byte
    burst records (the per-burst cost has a floor even for photon-less bursts)."""
    ph = np.asarray(ph_count, dtype=np.int64)
    B = ph.shape[0]
    if world < 1:
        raise ValueError("world must be >= 1")
    cost = np.cumsum(ph + 4)
    total = int(cost[-1]) if B else 0
    cuts = [0]
    for r in range(1, world):
        target = total * r / world
        cuts.append(int(np.searchsorted(cost, target, side="left")) if B else 0)
    cuts.append(B)
    for i in range(1, len(cuts)):
        cuts[i] = max(cuts[i], cuts[i - 1])
    return [(cuts[i], cuts[i + 1]) for i in range(world)]


def split_population(P: int, world: int):
    """-> list of (p0, p1): contiguous shares of a GA population (no communication needed)."""
    base, extra = divmod(P, world)
    out, p = [], 0
    for r in range(world):
        n = base + (1 if r < extra else 0)
        out.append((p, p + n))
        p += n
    return out


def expected_jumps(n: int, args, T: float) -> float:
    """Expected number of state changes within T of the walk mc_kernel takes under parameter vector
    `args` (layout of mc::set_params: E | K off-diagonal, row-major | v).  The walk starts in state 0:
    the reference compares the UNNORMALISED inflow rates with a uniform in (0, 1]
    (drawDisIdx, /root/reference/src/gen_rand.cuh:35-50 called from mc.cu:110), so the first running sum
    — any rate above 1/s — always wins.  E[N_T] = q^T (int_0^T e^{Gs} ds) e_0, G the generator with
    K(row, col) = rate col -> row, q the exit rates."""
    a = np.asarray(args, dtype=np.float64)
    rates = iter(a[n:n + n * n - n])
    K = np.zeros((n, n))
    for row in range(n):
        for col in range(n):
            if row != col:
                K[row, col] = next(rates)
    q = K.sum(axis=0)
    G = K - np.diag(q)
    try:
        lam, V = np.linalg.eig(G)
        with np.errstate(divide="ignore", invalid="ignore"):
            small = np.abs(lam * T) < 1e-9
            integ = np.where(small, T, np.expm1(lam * T) / np.where(small, 1.0, lam))
        c = np.linalg.solve(V, np.eye(n)[:, 0].astype(complex))
        val = float(np.real(q @ (V @ (integ * c))))
        if np.isfinite(val) and val >= 0.0:
            return val
    except np.linalg.LinAlgError:
        pass
    return float(q[0] * T)


def vector_cost(n: int, args, mean_T: float = 2.0e-3):
    """Relative cost of one burst-sample under parameter vector `args`: one dwell window plus one per
    expected state change within a burst of mean duration.  Calibrated on a B200 with
    tools/calibrate_vector_cost.py (32 vectors of the config-4 population, 1e5 bursts x 16:
    ms = 0.1125 + 0.0491 E[jumps], rms residual 8 %; profiles/r03f_vector_cost.txt) — round 1's model
    took the stationary exit flux, which misses that every sample starts in state 0."""
    return 1.0 + 0.436 * expected_jumps(n, args, mean_T)


def deal_population(costs, world: int):
    """-> one index array per rank: the vectors sorted by descending cost and dealt in snake order
    (0..w-1, w-1..0, ...), so that every rank's share costs about the same although the vectors do
    not (a vector with fast interconversion walks many more dwell segments per burst)."""
    order = np.argsort(-np.asarray(costs, dtype=np.float64), kind="stable")
    shares = [[] for _ in range(world)]
    for pos, idx in enumerate(order):
        rnd, k = divmod(pos, world)
        shares[k if rnd % 2 == 0 else world - 1 - k].append(int(idx))
    return [np.array(sorted(s), dtype=np.int64) for s in shares]
