"""The C++ host side (gsmfretda_b200/host): `class mc` with the reference's signatures over the
C ABI, the gpuWorker state machine, and local_node (parameter threads + gpuWorker thread)."""
import os
import subprocess

import numpy as np
import pytest

from gsmfretda_b200 import synth
from helpers import NB, LEFT_PAD

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "gsmfretda_b200", "host")


def build():
    subprocess.run(["make", "-s", "-C", HOST], check=True)
    return os.path.join(HOST, "local_node")


def test_shim_compiles_and_links():
    exe = build()
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 2 and "usage" in r.stderr


def test_shim_declares_the_reference_members():
    src = open(os.path.join(HOST, "mc.hpp")).read()
    for member in ("void set_gpuid()", "cudaStream_t* getStreams(int* numStream", "bool streamQuery(int sid)",
                   "int set_nstates(int n, int sid)", "bool set_params(int n, int sid, std::vector<float>& args)",
                   "int setBurstBd(int cstart, int cstop, int sid)", "void set_reSampleTimes(int t)",
                   "void init_randstate(int N, int sid)", "void run_kernel(int N, int sid)",
                   "void set_params_buff(int oldS_n, int N_sid, int sid)", "void init_data_gpu(",
                   "bool get_max_conversion_capacity(int max_stateNum = 8)", "std::atomic_int workerNum",
                   "retype** hmcE", "int* s_n", "int sz_burst", "int reSampleTimes", "char gpuuuid[33]"):
        assert member in src, member


REF_SRC = "/root/reference/src"
DROPIN = os.path.join(HOST, "dropin_node")


@pytest.mark.skipif(not os.path.exists(os.path.join(REF_SRC, "gpuWorker.cpp")), reason="the reference checkout lives on the build box only")
def test_reference_gpuworker_compiles_unchanged_against_the_shim():
    """Drop-in proof: /root/reference/src/gpuWorker.cpp + gpuWorker.hpp, byte for
    byte as they lie in the reference, compile against host/mc.hpp (with compat/args.pb.h and
    compat/tools.hpp for the two other headers they include) and link with libgsmfret.so into
    local_node's main.  The binary travels to the GPU box, where the test below runs it."""
    build()
    if os.path.exists(DROPIN):
        os.unlink(DROPIN)
    r = subprocess.run(["make", "-s", "-C", HOST, "dropin"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    assert os.path.exists(DROPIN)
    # it is the reference's class that was linked: its run() prints this line with -G (debug gpu)
    syms = subprocess.run(["nm", "-C", DROPIN], capture_output=True, text=True).stdout
    assert "gpuWorker::run(int)" in syms and "gpuWorker::gpuWorker(" in syms
    strings = subprocess.run(["strings", DROPIN], capture_output=True, text=True).stdout
    assert "gpu dataready ==3" in strings      # a literal of /root/reference/src/gpuWorker.cpp:50
    r = subprocess.run([DROPIN], capture_output=True, text=True)
    assert r.returncode == 2 and "usage" in r.stderr


def _check_node_output(oracle, exe, tmp_path, streams, coalesce=0, P=10, batch=1):
    w = synth.workload(4, n_bursts=1200, R=8, P=P)
    b = w.bursts()
    synth.save_bin(b, str(tmp_path / "b.bin"))
    with open(tmp_path / "p.txt", "w") as f:
        for a in w.args:
            f.write("2 " + " ".join(repr(float(x)) for x in a) + "\n")
        f.write("3 " + " ".join(repr(float(x)) for x in synth.workload(3).args[0]) + "\n")   # n changes mid-run
    r = subprocess.run([exe, str(tmp_path / "b.bin"), str(tmp_path / "p.txt"), str(tmp_path / "o.txt"), str(streams), "8", "42"]
                       + ([str(coalesce), "0", "16", "60", str(batch)] if (coalesce or batch > 1) else []),
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    rows = [l.split() for l in open(tmp_path / "o.txt") if l.strip()]
    assert len(rows) == P + 1
    ohist = oracle.observed_hist(b, NB)
    allargs = [(2, a) for a in w.args] + [(3, synth.workload(3).args[0])]
    for row in rows:
        ji, epoch, chi2 = int(row[0]), int(row[2]), float(row[3])
        hist = np.array([int(x) for x in row[4:]], np.uint64)
        n, a = allargs[ji]
        o = oracle.eval(b, n, a, 8, 42, epoch)
        oh = oracle.hist(o["mcE"], NB)
        assert np.array_equal(hist, oh), (ji, epoch, int(hist.sum()), int(oh.sum()))
        c = oracle.chi2(ohist, oh, NB, 8, LEFT_PAD, n)
        assert abs(chi2 - c) <= 2e-6 * abs(c)


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(DROPIN), reason="host/dropin_node is built where the reference checkout is (make -C host dropin)")
def test_reference_gpuworker_drives_the_library(oracle, tmp_path):
    """The reference's own gpuWorker::run (compiled unchanged, see above) in local_node's place:
    same results as the oracle, bit for bit, through set_nstates / set_params / setBurstBd /
    init_randstate / set_params_buff / run_kernel / streamQuery as the reference issues them."""
    _check_node_output(oracle, DROPIN, tmp_path, 3)


@pytest.mark.gpu
@pytest.mark.parametrize("streams,coalesce", [(8, 2), (16, 1), (3, 3)])
def test_local_node_coalesced_launches_match_oracle(oracle, tmp_path, streams, coalesce):
    """gpuWorker::set_coalesce: vectors that are ready together ride ONE launch on a carrier slot
    (gsm_run_coalesced), each with its own stream's draw epoch — the node's output is the oracle's,
    bit for bit, like the one-launch-per-stream node's (the vector of another model in the list is
    launched on its own)."""
    _check_node_output(oracle, build(), tmp_path, streams, coalesce=coalesce, P=60)


@pytest.mark.gpu
@pytest.mark.parametrize("streams,batch", [(3, 7), (2, 64)])
def test_local_node_batched_requests_match_oracle(oracle, tmp_path, streams, batch):
    """Every parameter thread asks for `batch` vectors at a time (the batched p_ga): one launch per
    request through mc::set_params_batch, vector p drawn with epoch + p; same output, bit for bit."""
    _check_node_output(oracle, build(), tmp_path, streams, batch=batch, P=60)


@pytest.mark.gpu
def test_local_node_matches_oracle(oracle, tmp_path):
    exe = build()
    w = synth.workload(4, n_bursts=1200, R=8, P=10)
    b = w.bursts()
    synth.save_bin(b, str(tmp_path / "b.bin"))
    with open(tmp_path / "p.txt", "w") as f:
        for a in w.args:
            f.write("2 " + " ".join(repr(float(x)) for x in a) + "\n")
        f.write("3 " + " ".join(repr(float(x)) for x in synth.workload(3).args[0]) + "\n")   # n changes mid-run
    r = subprocess.run([exe, str(tmp_path / "b.bin"), str(tmp_path / "p.txt"), str(tmp_path / "o.txt"), "3", "8", "42"],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    assert "param-evals/s" in r.stdout
    rows = [l.split() for l in open(tmp_path / "o.txt") if l.strip()]
    assert len(rows) == 11
    ohist = oracle.observed_hist(b, NB)
    allargs = [(2, a) for a in w.args] + [(3, synth.workload(3).args[0])]
    for row in rows:
        ji, epoch, chi2 = int(row[0]), int(row[2]), float(row[3])
        hist = np.array([int(x) for x in row[4:]], np.uint64)
        n, a = allargs[ji]
        o = oracle.eval(b, n, a, 8, 42, epoch)
        oh = oracle.hist(o["mcE"], NB)
        assert np.array_equal(hist, oh), (ji, epoch, int(hist.sum()), int(oh.sum()))
        c = oracle.chi2(ohist, oh, NB, 8, LEFT_PAD, n)
        assert abs(chi2 - c) <= 2e-6 * abs(c)


def test_split_node_plans_like_sharding_py(tmp_path):
    """The C++ burst-range planner of split_node equals gsmfretda_b200.sharding.split_bursts."""
    from gsmfretda_b200 import sharding
    build()
    exe = os.path.join(HOST, "split_node")
    for cfg, nb in ((2, 3000), (5, 777), (1, 5)):
        b = synth.workload(cfg, n_bursts=nb).bursts()
        synth.save_bin(b, str(tmp_path / "b.bin"))
        for parts in (1, 2, 3, 8):
            r = subprocess.run([exe, "--plan", str(tmp_path / "b.bin"), str(parts)], capture_output=True, text=True, timeout=60)
            assert r.returncode == 0, r.stderr
            got = [tuple(int(x) for x in l.split()) for l in r.stdout.splitlines() if l.strip()]
            assert got == sharding.split_bursts(b.ph_count, parts), (cfg, nb, parts)


@pytest.mark.gpu
@pytest.mark.parametrize("parts,over", [(1, False), (2, True), (5, True), (5, False)])
def test_split_node_matches_oracle(oracle, gsm, tmp_path, parts, over):
    """One dataset cut into `parts` burst ranges behind ONE multi-GPU `mc` (host/mc.hpp), histogram
    exchange fused into the evaluation kernel, driven from C++: the global histogram and chi-square
    equal the oracle's over the whole dataset, whatever the number of parts.  Parts beyond the
    number of GPUs are merged unless --oversubscribe keeps them apart on shared GPUs."""
    build()
    ngpu = gsm.load().gsm_device_count()
    used = parts if (over or parts <= ngpu) else ngpu
    exe = os.path.join(HOST, "split_node")
    w = synth.workload(5, n_bursts=1500, R=4)
    b = w.bursts()
    synth.save_bin(b, str(tmp_path / "b.bin"))
    vecs = [(w.n_states, w.args[0]), (2, synth.workload(2).args[0]), (3, synth.workload(3).args[0])]
    with open(tmp_path / "p.txt", "w") as f:
        for n, a in vecs:
            f.write(f"{n} " + " ".join(repr(float(x)) for x in a) + "\n")
    r = subprocess.run([exe, str(tmp_path / "b.bin"), str(tmp_path / "p.txt"), str(tmp_path / "o.txt"), str(parts), "4", "42"]
                       + (["--oversubscribe"] if over else []), capture_output=True, text=True, timeout=180)
    assert r.returncode == 0, (r.stdout, r.stderr)
    assert f"split into {used} part(s) ({parts} asked)" in r.stdout
    rows = [l.split() for l in open(tmp_path / "o.txt") if l.strip()]
    assert len(rows) == len(vecs)
    ohist = oracle.observed_hist(b, NB)
    for row in rows:
        ji, epoch, chi2 = int(row[0]), int(row[2]), float(row[3])
        hist = np.array([int(x) for x in row[4:]], np.uint64)
        n, a = vecs[ji]
        oh = oracle.hist(oracle.eval(b, n, a, 4, 42, epoch)["mcE"], NB)
        assert np.array_equal(hist, oh), (ji, int(hist.sum()), int(oh.sum()))
        c = oracle.chi2(ohist, oh, NB, 4, LEFT_PAD, n)
        assert abs(chi2 - c) <= 2e-6 * abs(c)
