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
