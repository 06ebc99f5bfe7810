"""Burst-split evaluation across ranks (host logic), world_size 2 and 3 over gloo on CPU:
each rank evaluates its contiguous burst range with global burst ids, the integer histograms
are all-reduced, and the sum equals the single-rank histogram bit for bit."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gsmfretda_b200 import sharding, synth
from helpers import NB, LEFT_PAD


def test_split_bursts_balances_photons():
    rng = np.random.default_rng(0)
    ph = rng.integers(0, 300, 10007)
    for world in (1, 2, 3, 8):
        parts = sharding.split_bursts(ph, world)
        assert parts[0][0] == 0 and parts[-1][1] == ph.size
        assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
        loads = [ph[a:b].sum() for a, b in parts]
        assert max(loads) - min(loads) <= 0.02 * sum(loads) / world + 600
    assert sharding.split_bursts(np.zeros(0, np.int32), 2) == [(0, 0), (0, 0)]
    assert sharding.split_population(1024, 8) == [(i * 128, (i + 1) * 128) for i in range(8)]
    assert [b - a for a, b in sharding.split_population(10, 4)] == [3, 3, 2, 2]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root); sys.path.insert(0, os.path.join(root, "oracle"))
    import oracle_py
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    O = oracle_py.Oracle()
    w = synth.workload(5, n_bursts=900, R=4)
    full = w.bursts()
    b0, b1 = sharding.split_bursts(full.ph_count, world)[rank]
    mine = full.slice(b0, b1)                       # what gsm_init_data(..., g_offset=b0) receives
    o = O.eval(mine, w.n_states, w.args[0], w.R, 77, 5, g_offset=b0, nthreads=2)
    h = torch.from_numpy(O.hist(o["mcE"], NB).astype(np.int64))
    oh = torch.from_numpy(O.observed_hist(mine, NB).astype(np.float32))
    dist.all_reduce(h)
    dist.all_reduce(oh)
    chi2 = O.chi2(oh.numpy(), h.numpy().astype(np.uint64), NB, w.R, LEFT_PAD, w.n_states)
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), h=h.numpy(), chi2=chi2)
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_burst_split_histogram_is_rank_count_invariant(oracle, tmp_path, world):
    port = _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    w = synth.workload(5, n_bursts=900, R=4)
    full = w.bursts()
    o = oracle.eval(full, w.n_states, w.args[0], w.R, 77, 5)
    h = oracle.hist(o["mcE"], NB)
    c = oracle.chi2(oracle.observed_hist(full, NB), h, NB, w.R, LEFT_PAD, w.n_states)
    for r in range(world):
        d = np.load(tmp_path / f"r{r}.npz")
        assert np.array_equal(d["h"].astype(np.uint64), h)
        assert np.float32(d["chi2"]) == np.float32(c)
