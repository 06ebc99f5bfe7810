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
