"""The worker node end to end: `gSMFRETda` reads an HDF5 burst file, talks
to a parameter server over nanomsg REQ/REP + protobuf, evaluates every vector on the GPU and
sends chi-square (+ the model histogram when it beats the best value) back.  The server here is
tests/sp_server.py (raw sockets + Google's protobuf runtime); the numbers are checked against the
oracle, bit for bit, in replay mode with draws keyed by the parameter index."""
import os
import subprocess

import numpy as np
import pytest

from gsmfretda_b200 import synth
from helpers import NB, LEFT_PAD
from sp_server import ParamServer

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "gsmfretda_b200", "host")
H5 = os.path.join(ROOT, "tests", "golden", "h5")
NODE = os.path.join(HOST, "gSMFRETda")
CALL = os.path.join(HOST, "call_gSMFRETda")
R, SEED = 8, 77

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def world(oracle):
    subprocess.run(["make", "-s", "-C", HOST], check=True)
    w = synth.workload(1, n_bursts=300, R=4)          # the bursts of tests/golden/h5/from_bin.h5
    b = w.bursts()
    rng = np.random.default_rng(3)
    vecs = [[float(np.float32(x)) for x in (rng.uniform(0.1, 0.45), rng.uniform(0.55, 0.9), rng.uniform(50, 2000),
                                             rng.uniform(50, 2000), rng.uniform(1, 4), rng.uniform(1, 4))]
            for _ in range(9)]
    ohist = oracle.observed_hist(b, NB)
    want = []
    for i, v in enumerate(vecs):
        o = oracle.eval(b, 2, np.array(v, np.float32), R, SEED, i + 1)     # --epoch-from-idx: epoch = idx + 1
        h = oracle.hist(o["mcE"], NB)
        want.append((h, oracle.chi2(ohist, h, NB, R, LEFT_PAD, 2)))
    return dict(b=b, vecs=vecs, ohist=ohist, want=want)


def run_node(srv, infile, *extra, exe=NODE, streams=3):
    cmd = [exe, "-u", srv.url, "-i", infile, "-s", str(streams), "--mode", "replay", "--epoch-from-idx",
           "--resample-times", str(R), "--seed", str(SEED), *extra]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=180)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])
    return r


def check(srv, world, bestfv):
    assert sorted(srv.results) == list(range(len(world["vecs"])))
    assert srv.caps and set(srv.caps) == {300}                      # 'c': cap = number of bursts
    nodes = [n for n in srv.nodes.values() if n["ohist"] is not None]
    assert nodes, "every stream sends the observed histogram with its first 'p'"
    for n in nodes:
        assert n["histNum"] == NB
        assert np.array_equal(np.array(n["ohist"], np.float32), world["ohist"][:NB].astype(np.float32))
    for i, (h, c) in enumerate(world["want"]):
        got = srv.results[i]
        assert np.float32(got["e"]) == np.float32(c), (i, got["e"], c)
        assert got["hist"] == (np.float32(c) < np.float32(bestfv))
        if got["hist"]:
            assert np.array_equal(np.array(got["shist"], np.float32), (h[:NB].astype(np.float32) / np.float32(R)))
        else:
            assert got["shist"] == []
        # node id: 32 hex digits of the GPU uuid - pid - thread id - gpu - stream (tools.cpp:155-167)
        f = got["idx"].split("-")
        assert len(f) == 5 and len(f[0]) == 32 and f[3] == "0" and 0 <= int(f[4]) < 3


def test_node_from_hdf5_file_matches_the_oracle(world):
    best = float(np.median([c for _, c in world["want"]]))           # about half the results carry shist
    srv = ParamServer(2, world["vecs"], bestfv=best, idle_first=2)
    try:
        r = run_node(srv, os.path.join(H5, "from_bin.h5"))
        assert "loaded." in r.stdout and "streamNum: 3" in r.stdout and "gpu thread joined" in r.stdout
        check(srv, world, best)
    finally:
        srv.close()


@pytest.mark.parametrize("extra", [(), ("--host-hist",)], ids=["fused", "host-hist"])
def test_batched_p_ga_is_one_launch_per_batch_and_the_same_numbers(world, extra):
    """Batch extension: the server packs 4 vectors into one p_ga; the node evaluates
    them with one gsm_run_batch launch and answers with one `res` carrying es[] and best.  Every
    chi-square equals the unbatched one bit for bit (draws keyed by the vector index); the
    histogram attached is the best vector's of each batch."""
    best = float(np.median([c for _, c in world["want"]]))
    srv = ParamServer(2, world["vecs"], bestfv=best, batch=4)
    try:
        run_node(srv, os.path.join(H5, "from_bin.h5"), *extra, streams=2)
        assert sorted(srv.results) == list(range(len(world["vecs"])))
        assert len(srv.raw) == 3                                      # 9 vectors in batches of 4, 4, 1
        for first in (0, 4, 8):
            members = list(range(first, min(first + 4, 9)))
            cs = [np.float32(world["want"][i][1]) for i in members]
            k = int(np.argmin(cs))
            for j, i in enumerate(members):
                got = srv.results[i]
                assert np.float32(got["e"]) == cs[j], (i, got["e"], cs[j])
                attach = j == k and cs[k] < np.float32(best)
                assert got["hist"] == attach
                if attach:
                    assert np.array_equal(np.array(got["shist"], np.float32),
                                          world["want"][i][0][:NB].astype(np.float32) / np.float32(R))
                else:
                    assert got["shist"] == []
    finally:
        srv.close()


@pytest.mark.parametrize("batch", [1, 4])
def test_one_node_spanning_gpus_matches_the_oracle(world, gsm, batch):
    """`gSMFRETda -g a,b`: ONE node, one set of REQ clients, the dataset split over the listed GPUs
    inside the process (multi-GPU `mc`, fused peer exchange) — what replaces one node per GPU
    (/root/reference/src/call_main.cpp:195-227) when it is the dataset that is large.  Same wire
    conversation, same numbers as the oracle over the whole dataset, single vectors and batches."""
    ndev = gsm.load().gsm_device_count()
    gpus, extra = ("0,1", ()) if ndev >= 2 else ("0,0", ("--oversubscribe",))
    best = float(np.median([c for _, c in world["want"]]))
    srv = ParamServer(2, world["vecs"], bestfv=best, batch=batch)
    try:
        r = run_node(srv, os.path.join(H5, "from_bin.h5"), "-g", gpus, *extra, streams=2)
        assert "GPUs: " + gpus.replace(",", " ") in r.stdout
        assert sorted(srv.results) == list(range(len(world["vecs"])))
        for i, (h, c) in enumerate(world["want"]):
            assert np.float32(srv.results[i]["e"]) == np.float32(c), (i, srv.results[i]["e"], c)
            if srv.results[i]["hist"]:
                assert np.array_equal(np.array(srv.results[i]["shist"], np.float32), h[:NB].astype(np.float32) / np.float32(R))
        assert any(v["hist"] for v in srv.results.values())
    finally:
        srv.close()


def test_a_refused_vector_is_reported_not_evaluated(world):
    """A vector the library refuses (here: the server announces 1 state, genMatK's empty rate list,
    eigenhelper.cpp:5-8) launches nothing: chi-square +inf, no histogram — never the numbers of the
    previous vector of that stream."""
    srv = ParamServer(1, [[0.5, 2.0]] * 3, bestfv=3.0e32)
    try:
        run_node(srv, os.path.join(H5, "from_bin.h5"), streams=1)
        assert sorted(srv.results) == [0, 1, 2]
        for i in range(3):
            assert srv.results[i]["e"] == float("inf") and srv.results[i]["hist"] is False and srv.results[i]["shist"] == []
    finally:
        srv.close()


def test_host_histogram_path_sends_the_same_numbers(world):
    # --host-hist: hmcE copied back, histogram and chi-square on the host with the reference's
    # arithmetic (streamWorker.cpp:186-206) -> same bytes as the fused kernel's
    srv = ParamServer(2, world["vecs"], bestfv=3.0e32)
    try:
        run_node(srv, os.path.join(H5, "cfg1_300.bin"), "--host-hist")
        check(srv, world, 3.0e32)
    finally:
        srv.close()


def test_launcher_starts_one_node_per_gpu(world, tmp_path):
    srv = ParamServer(2, world["vecs"], bestfv=0.0)
    try:
        r = run_node(srv, os.path.join(H5, "from_bin.h5"), "-g", "0", "-I", str(tmp_path / "nodes.pid"), exe=CALL)
        assert "exit code:0" in r.stdout
        check(srv, world, 0.0)
        pids = open(tmp_path / "nodes.pid").read().split()
        assert len(pids) == 1                                       # the node appended its pid (main.cpp:53-64)
    finally:
        srv.close()


def test_one_vector_from_the_command_line(world, gsm):
    v = world["vecs"][4]
    r = subprocess.run([NODE, "-i", os.path.join(H5, "from_bin.h5"), "-s", "1", "--mode", "replay", "--resample-times",
                        str(R), "--seed", str(SEED), "-p", ",".join(repr(x) for x in v)],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    chi2 = float([l for l in r.stdout.split("\n") if l.startswith("chi2:")][0].split()[1])
    with gsm.Mc(0, 1) as mc:
        mc.set_reSampleTimes(R)
        mc.set_seed(SEED)
        mc.set_hist(NB, LEFT_PAD)
        mc.init_bursts(world["b"])
        c, _ = mc.evaluate(2, np.array(v, np.float32), sid=0)         # first evaluation of a fresh context
    assert np.float32(chi2) == np.float32(c)
    bad = subprocess.run([NODE, "-i", os.path.join(H5, "from_bin.h5"), "-p", "0.3,0.7,1"], capture_output=True, text=True)
    assert bad.returncode != 0 and "n*n+n" in bad.stdout


def test_normal_mode_node_is_statistically_the_same(world):
    srv = ParamServer(2, world["vecs"][:4], bestfv=3.0e32)
    try:
        cmd = [NODE, "-u", srv.url, "-i", os.path.join(H5, "from_bin.h5"), "-s", "2", "--resample-times", "64"]
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=180)
        assert r.returncode == 0, r.stderr
        for i in range(4):
            sh = np.array(srv.results[i]["shist"])
            # model histogram / R: the bursts, minus the samples with E = 1 exactly (overflow bin, Q9)
            assert 250 < sh.sum() <= 300 + 1e-3
            h, _ = world["want"][i]
            # 64 repeats against 8: the normalised histograms agree within Monte-Carlo noise
            assert np.abs(sh / sh.sum() - h[:NB] / h[:NB].sum()).sum() < 0.5
    finally:
        srv.close()


def test_node_in_reference_mode_equals_the_reference(world, ref, oracle):
    """gSMFRETda --mode reference from an HDF5 file: the chi-square printed for one vector is the
    one streamWorker would compute from the reference's own mc_kernel samples (host build of the
    reference sources on cuRAND's Sobol draws = the reference's GPU run, test_reference_gpu_golden.py)."""
    v = world["vecs"][2]
    r = subprocess.run([NODE, "-i", os.path.join(H5, "from_bin.h5"), "-s", "1", "--mode", "reference", "--resample-times",
                        str(R), "-p", ",".join(repr(x) for x in v)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    chi2 = float([l for l in r.stdout.split("\n") if l.startswith("chi2:")][0].split()[1])
    e = ref.mc_kernel_sobol(world["b"], 2, np.array(v, np.float32), R, evals=1)[0]
    want = oracle.chi2(world["ohist"], oracle.hist(e, NB), NB, R, LEFT_PAD, 2)
    assert np.float32(chi2) == np.float32(want), (chi2, want)


def test_coalesce_in_reference_mode_falls_back_to_one_launch_per_stream(world):
    """--coalesce with --mode reference: reference draws are one vector per launch, so the node says that
    coalescing is off and serves every vector with per-stream launches instead of dying on the library's
    refusal (GSM_ESTATE) in the middle of an optimisation."""
    srv = ParamServer(2, world["vecs"], bestfv=3.0e32)
    try:
        cmd = [NODE, "-u", srv.url, "-i", os.path.join(H5, "from_bin.h5"), "-s", "3", "--mode", "reference",
               "--coalesce", "2", "--resample-times", str(R)]
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=180)
        assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])
        assert "--coalesce is off: reference mode" in r.stdout
        assert sorted(srv.results) == list(range(len(world["vecs"])))
        for i in range(len(world["vecs"])):
            assert np.isfinite(srv.results[i]["e"]) and srv.results[i]["e"] > 0
    finally:
        srv.close()
