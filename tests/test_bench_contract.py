"""bench.py's line contract, as far as a CPU box can check it: the reference arm runs here (it
times the reference's CPU path), prints exactly one JSON line with the keys the driver reads, and
the non-zero ranks of a torchrun launch leave without work."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(env_extra=None, *args):
    env = dict(os.environ)
    env.update(env_extra or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                          timeout=300, env=env, cwd=ROOT)


def test_reference_arm_prints_one_json_line():
    r = run({"OMP_NUM_THREADS": "1"}, "--impl", "reference", "--steps", "2", "--warmup", "1", "--bursts", "4000")
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "burst-samples/s" and d["unit"] == "burst-samples/s"
    assert d["higher_is_better"] is True and d["steps"] == 2 and d["warmup"] == 1 and d["n_gpus"] == 1
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["gpu_launches"] == 0 and d["vs_baseline"] is None
    assert d["config"]["workload"].startswith("cfg2")
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["value"] == d["value"] and cb["sample"]
    assert cb["cores"] >= 1 and (cb["cores"] == len(os.sched_getaffinity(0)))      # not OMP_NUM_THREADS
    assert d["e2e"] == {"value": d["value"], "unit": "burst-samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_exit_without_work():
    r = run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}, "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0")
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_our_arm_refuses_to_run_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        return
    r = run(None, "--steps", "1", "--warmup", "0", "--no-cpu", "--bursts", "100")
    assert r.returncode != 0 and "no CPU fallback" in (r.stderr + r.stdout)


def _bench_module():
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_both_arms_label_the_workload_with_one_function():
    """`config` of the reference arm and of the GPU arm come from the same function with the same
    inputs, so the driver's same-config check cannot trip over wording: repeats = what ran (100, as
    passes of 64 + 32 + 4 on the reference, which keeps a power of two per pass)."""
    b = _bench_module()
    assert b.pow2_passes(100) == [64, 32, 4] and b.pow2_passes(16) == [16] and b.pow2_passes(7) == [4, 2, 1]
    r = run({"OMP_NUM_THREADS": "1"}, "--impl", "reference", "--steps", "1", "--warmup", "0")
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads(r.stdout.strip().splitlines()[-1])
    from gsmfretda_b200 import synth
    w = synth.workload(2)
    bursts = w.bursts()
    ours = b.workload_config(w, bursts.n_bursts, bursts.n_photons, 1, 1, b.sharding_text("split", 1, "peer"), bursts.n_bursts)
    assert d["config"] == ours
    assert d["config"]["repeats"] == 100 and "64+32+4" in d["cpu_baseline"]["sample"]
    assert d["scaling"] == "weak" and d["metric"] == "burst-samples/s"


def test_population_is_dealt_in_shares_of_equal_cost():
    from gsmfretda_b200 import sharding, synth
    w = synth.workload(4)
    cost = [sharding.vector_cost(2, a) for a in w.args]
    for world in (2, 3, 8):
        shares = sharding.deal_population(cost, world)
        assert sorted(int(i) for s in shares for i in s) == list(range(1024))
        tot = [sum(cost[i] for i in s) for s in shares]
        assert max(tot) / min(tot) < 1.01, (world, tot)
        assert max(len(s) for s in shares) - min(len(s) for s in shares) <= 1
    # fast interconversion costs more than slow
    slow = synth.pack_args([0.3, 0.7], [10, 10], [2, 2])
    fast = synth.pack_args([0.3, 0.7], [5000, 5000], [2, 2])
    assert sharding.vector_cost(2, fast) > 3 * sharding.vector_cost(2, slow)


def test_cost_model_counts_the_jumps_the_walk_takes(oracle):
    """sharding.expected_jumps against scipy's matrix exponential, and against the walk itself: the
    oracle's dwell-window count per sample is 1 + the state changes within the burst, starting — as the
    reference does — in state 0 whatever the stationary distribution says."""
    import numpy as np
    from scipy.linalg import expm
    from gsmfretda_b200 import sharding, synth

    def by_expm(n, args, T):
        a = np.asarray(args, dtype=np.float64)
        it = iter(a[n:n + n * n - n])
        K = np.zeros((n, n))
        for row in range(n):
            for col in range(n):
                if row != col:
                    K[row, col] = next(it)
        q = K.sum(axis=0)
        M = np.zeros((2 * n, 2 * n))
        M[:n, :n] = K - np.diag(q)
        M[:n, n:] = np.eye(n)
        return float(q @ expm(M * T)[:n, n:][:, 0])

    for cfg in (2, 3, 4, 5):
        w = synth.workload(cfg)
        for a in w.args[:20]:
            for T in (2e-4, 2e-3, 3e-2):
                assert abs(sharding.expected_jumps(w.n_states, a, T) - by_expm(w.n_states, a, T)) < 1e-9 * (1 + by_expm(w.n_states, a, T))
    w = synth.workload(4)
    b = w.bursts(n_bursts=1500)
    for i in (158, 770, 603, 1011):   # slow/slow, slow/fast (stationary flux ~0 but always one jump), fast/slow, fast/fast
        steps = oracle.eval(b, 2, w.args[i], 8, 3, 0)["n_steps"]
        want = 1.0 + float(np.mean([sharding.expected_jumps(2, w.args[i], float(T)) for T in b.T]))
        assert abs(float(np.mean(steps)) - want) < 0.04 * want, (i, float(np.mean(steps)), want)
