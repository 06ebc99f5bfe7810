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


def vector_cost(n: int, args, mean_T: float = 2.0e-3, mean_photons: float = 60.0):
    """Rough relative cost of one burst-sample under parameter vector `args` (layout of
    mc::set_params: E | K off-diagonal, row-major | v): photons walked plus dwell segments, the
    latter from the stationary exit flux of K times the mean burst duration.  Only the ORDER of the
    costs matters to deal_population."""
    a = np.asarray(args, dtype=np.float64)
    k = a[n:n + n * n - n]
    Q = np.zeros((n, n))
    it = iter(k)
    for i in range(n):
        for j in range(n):
            if i != j:
                Q[i, j] = next(it)
    q = Q.sum(axis=1)
    A = (Q - np.diag(q)).T
    A[-1, :] = 1.0
    rhs = np.zeros(n)
    rhs[-1] = 1.0
    try:
        pi = np.linalg.solve(A, rhs)
    except np.linalg.LinAlgError:
        pi = np.full(n, 1.0 / n)
    flux = float(np.dot(np.clip(pi, 0.0, 1.0), q))
    return 6.0 * mean_photons + (80.0 + 6.0 * n) * (1.0 + flux * mean_T)


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
