"""Fused cross-GPU histogram exchange (gsm_peer_export / gsm_peer_connect, SURVEY.md §8e burst split).

Every rank holds a contiguous burst range; the last block of each rank's evaluation launch pushes
its partial histogram into every rank's inbox, so each rank reports the global histogram and
chi-square from ONE launch.  Integer bins make the result independent of the rank count: the
checks are bit-exact against the oracle over the whole dataset and against a single context.

The contexts of the in-process tests share device 0 (the round-end box has one GPU): the
exchange then runs through ordinary device pointers (conftest.py raises
CUDA_DEVICE_MAX_CONNECTIONS so that every rank's stream has its own hardware queue); the two-process test maps the inboxes with
CUDA IPC exactly as a one-process-per-GPU job does.
"""
import os
import subprocess
import sys

import numpy as np
import pytest

from gsmfretda_b200 import sharding, synth
from helpers import LEFT_PAD, NB

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _connected(gsm, b, world, R, seed, max_P=1, streams=1, mode=0):
    parts = sharding.split_bursts(b.ph_count, world)
    mcs = []
    ndev = max(1, gsm.load().gsm_device_count())
    for r, (b0, b1) in enumerate(parts):
        # rank r on GPU r % #GPUs: real peer memory over NVLink where the box has several GPUs; ranks
        # that have to share a GPU say so (GSM_PEER_SHARED_DEVICE; conftest gives them hardware queues)
        mc = gsm.Mc(r % ndev, streams, peer_shared_device=world > ndev)
        mc.set_reSampleTimes(R); mc.set_seed(seed); mc.set_hist(NB, LEFT_PAD); mc.set_mode(mode)
        mc.init_bursts(b.slice(b0, b1), g_offset=b0)
        mcs.append(mc)
    oh = sum(mc.observed_hist() for mc in mcs)
    handles = [mc.peer_export(world, max_P) for mc in mcs]
    for r, mc in enumerate(mcs):
        mc.set_observed_hist(oh)
        mc.peer_connect(handles, r, world)
        mc.peer_set_timeout(30000)
    return mcs, parts


@pytest.mark.parametrize("cfg,world", [(2, 2), (3, 3), (5, 4), (1, 8)])
def test_peer_exchange_matches_oracle(gsm, oracle, cfg, world):
    w = synth.workload(cfg, n_bursts=3000, R=8)
    b = w.bursts()
    n, R = w.n_states, w.R
    o = oracle.eval(b, n, w.args[0], R, 31, 4)
    oh = oracle.hist(o["mcE"], NB)
    oc = oracle.chi2(oracle.observed_hist(b, NB), oh, NB, R, LEFT_PAD, n)
    mcs, parts = _connected(gsm, b, world, R, 31)
    try:
        for rep in range(3):          # the parity of the inbox alternates; epoch pinned: same bits
            for mc, (b0, b1) in zip(mcs, parts):
                mc.set_nstates(n, 0); mc.set_params(n, 0, w.args[0]); mc.set_epoch(0, 4)
                mc.run_kernel(b1 - b0, 0)
            for mc in mcs:
                mc.sync(0)
                assert mc.flags(0) == 0
                assert np.array_equal(mc.hist(0, 0), oh)
                assert np.float32(mc.chi2(0)) == np.float32(oc)
        launches = [mc.launch_count() for mc in mcs]
        for mc, (b0, b1) in zip(mcs, parts):
            mc.set_epoch(0, 5); mc.run_kernel(b1 - b0, 0)
        for mc in mcs:
            mc.sync(0)
        assert [mc.launch_count() - l for mc, l in zip(mcs, launches)] == [1] * world  # one launch each
    finally:
        for mc in mcs:
            mc.close()


def test_peer_exchange_batch_and_streams(gsm):
    """P vectors per launch on two stream slots: equal to one context over all the bursts."""
    w = synth.workload(4, n_bursts=2000, R=4)
    b = w.bursts()
    n, R, P = w.n_states, w.R, 6
    args = np.ascontiguousarray(w.args[:P])
    with gsm.Mc(0, 2) as one:
        one.set_reSampleTimes(R); one.set_seed(9); one.set_hist(NB, LEFT_PAD)
        one.init_bursts(b)
        want = {}
        for sid in (0, 1):
            one.set_epoch(sid, 100 * sid + 7)
            one.run_batch(b.n_bursts, sid, n, args)
            one.sync(sid)
            want[sid] = (one.chi2v(sid).copy(), [one.hist(sid, p).copy() for p in range(P)])
    mcs, parts = _connected(gsm, b, 2, R, 9, max_P=P, streams=2)
    try:
        for sid in (0, 1):
            for mc, (b0, b1) in zip(mcs, parts):
                mc.set_epoch(sid, 100 * sid + 7)
                mc.run_batch(b1 - b0, sid, n, args)
        for sid in (0, 1):
            for mc in mcs:
                mc.sync(sid)
                assert mc.flags(sid) == 0
                assert np.array_equal(mc.chi2v(sid).view(np.uint32), want[sid][0].view(np.uint32))
                for p in range(P):
                    assert np.array_equal(mc.hist(sid, p), want[sid][1][p])
        with pytest.raises(RuntimeError):     # more vectors than the exchange was sized for
            mcs[0].run_batch(10, 0, n, np.ascontiguousarray(w.args[:P + 1]))
    finally:
        for mc in mcs:
            mc.close()


def test_peer_timeout_is_reported_not_hung(gsm):
    """A rank whose peer never launches gives up after the timeout and flags the result."""
    w = synth.workload(2, n_bursts=400, R=4)
    b = w.bursts()
    mcs, parts = _connected(gsm, b, 2, w.R, 3)
    try:
        mcs[0].peer_set_timeout(300)
        mcs[0].set_nstates(2, 0); mcs[0].set_params(2, 0, w.args[0])
        mcs[0].run_kernel(parts[0][1] - parts[0][0], 0)
        mcs[0].sync(0)
        assert mcs[0].flags(0) & gsm.GSM_FLAG_PEER_TIMEOUT
    finally:
        for mc in mcs:
            mc.close()


def test_ranks_sharing_a_gpu_are_refused_without_the_flag(gsm):
    """The in-kernel wait needs all ranks running at once: two ranks on one GPU are only accepted
    from a caller that vouches for it, and never with more (ranks x slots) than hardware queues."""
    w = synth.workload(2, n_bursts=200, R=4)
    b = w.bursts()

    def pair(streams, **kw):
        mcs = [gsm.Mc(0, streams, **kw) for _ in range(2)]
        for r, mc in enumerate(mcs):
            mc.set_reSampleTimes(4); mc.init_bursts(b.slice(100 * r, 100 * r + 100), g_offset=100 * r)
        return mcs, [mc.peer_export(2, 1) for mc in mcs]

    mcs, hs = pair(1)
    try:
        with pytest.raises(RuntimeError, match="share GPU"):
            mcs[0].peer_connect(hs, 0, 2)
    finally:
        for mc in mcs:
            mc.close()
    queues = int(os.environ.get("CUDA_DEVICE_MAX_CONNECTIONS", "8"))
    mcs, hs = pair(queues, peer_shared_device=True)      # 2 ranks x `queues` slots: twice too many
    try:
        with pytest.raises(RuntimeError, match="hardware queues"):
            mcs[0].peer_connect(hs, 0, 2)
    finally:
        for mc in mcs:
            mc.close()
    mcs, hs = pair(1, peer_shared_device=True)
    try:
        for r, mc in enumerate(mcs):
            mc.peer_connect(hs, r, 2)
        with pytest.raises(RuntimeError, match="disconnect first"):   # peers have the block mapped
            mcs[0].peer_export(2, 1)
    finally:
        for mc in mcs:
            mc.close()


def test_peer_api_errors(gsm):
    w = synth.workload(2, n_bursts=100, R=4)
    b = w.bursts()
    with gsm.Mc(0, 1) as mc:
        mc.set_reSampleTimes(4); mc.init_bursts(b)
        with pytest.raises(RuntimeError):
            mc.peer_export(9, 1)                       # more than GSM_MAX_PEERS
        with pytest.raises(RuntimeError):
            mc.peer_connect(b"\0" * 256, 0, 2)         # connect before export
        h = mc.peer_export(2, 1)
        with pytest.raises(RuntimeError):
            mc.peer_connect(h + b"\0" * 128, 0, 2)     # a handle that is not an export
        mc.peer_disconnect()
        c, hist = mc.evaluate(2, w.args[0])            # back to the single-GPU path
        assert int(hist.sum()) == 100 * 4


_WORKER = r"""
import os, sys, numpy as np
sys.path.insert(0, {root!r})
from gsmfretda_b200 import capi, sharding, synth
rank, world, d = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
w = synth.workload(2, n_bursts=4000, R=8)
b = w.bursts()
b0, b1 = sharding.split_bursts(b.ph_count, world)[rank]
ndev = capi.load().gsm_device_count()
mc = capi.Mc(rank % ndev, 1, peer_shared_device=world > ndev)   # one GPU per rank where the box has them
mc.set_reSampleTimes(w.R); mc.set_seed(12); mc.set_hist(200, 2)
mc.init_bursts(b.slice(b0, b1), g_offset=b0)
open(os.path.join(d, f"h{{rank}}.tmp"), "wb").write(mc.peer_export(world, 1))
os.rename(os.path.join(d, f"h{{rank}}.tmp"), os.path.join(d, f"h{{rank}}"))
np.save(os.path.join(d, f"oh{{rank}}.tmp.npy"), mc.observed_hist())
os.rename(os.path.join(d, f"oh{{rank}}.tmp.npy"), os.path.join(d, f"oh{{rank}}.npy"))
import time
t0 = time.time()
while not all(os.path.exists(os.path.join(d, f"h{{r}}")) and os.path.exists(os.path.join(d, f"oh{{r}}.npy")) for r in range(world)):
    assert time.time() - t0 < 120
    time.sleep(0.02)
hs = b"".join(open(os.path.join(d, f"h{{r}}"), "rb").read() for r in range(world))
mc.set_observed_hist(sum(np.load(os.path.join(d, f"oh{{r}}.npy")) for r in range(world)))
mc.peer_connect(hs, rank, world)
mc.peer_set_timeout(60000)
out = []
for ep in (3, 4, 5):
    c, h = mc.evaluate(2, w.args[0], epoch=ep)
    assert mc.flags(0) == 0, mc.flags(0)
    out.append(np.concatenate([h.astype(np.uint64), np.array([np.float32(c).view(np.uint32)], np.uint64)]))
np.save(os.path.join(d, f"res{{rank}}.npy"), np.stack(out))
open(os.path.join(d, f"done{{rank}}"), "w").write("ok")
t0 = time.time()
while not all(os.path.exists(os.path.join(d, f"done{{r}}")) for r in range(world)):
    assert time.time() - t0 < 120
    time.sleep(0.02)
mc.close()
"""


def test_peer_exchange_two_processes_cuda_ipc(gsm, oracle, tmp_path):
    """One process per rank (one GPU each where the box has two, else both on device 0): inboxes
    mapped with cudaIpcOpenMemHandle."""
    world = 2
    script = tmp_path / "worker.py"
    script.write_text(_WORKER.format(root=ROOT))
    procs = [subprocess.Popen([sys.executable, str(script), str(r), str(world), str(tmp_path)],
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(world)]
    outs = []
    for p in procs:
        try:
            outs.append(p.communicate(timeout=300)[0])
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)
    w = synth.workload(2, n_bursts=4000, R=8)
    b = w.bursts()
    res = [np.load(tmp_path / f"res{r}.npy") for r in range(world)]
    assert np.array_equal(res[0], res[1])
    ohist = oracle.observed_hist(b, NB)
    for i, ep in enumerate((3, 4, 5)):
        o = oracle.eval(b, 2, w.args[0], w.R, 12, ep)
        oh = oracle.hist(o["mcE"], NB)
        oc = oracle.chi2(ohist, oh, NB, w.R, LEFT_PAD, 2)
        assert np.array_equal(res[0][i][:NB + 2], oh.astype(np.uint64))
        assert int(res[0][i][NB + 2]) == int(np.float32(oc).view(np.uint32))
