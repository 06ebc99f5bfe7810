#!/usr/bin/env python
"""bench.py — throughput of the Monte-Carlo PDA evaluation (BASELINE.json metric:
burst-samples/s) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--cfg 2]

A "step" is one evaluation of one parameter vector over the whole synthetic burst set
(config 2: 1e5 bursts x 100 repeats = 1e7 burst-samples per GPU).  At N > 1 (torchrun) every
rank holds its own 1e5-burst shard of one N x 1e5-burst dataset (weak scaling) and each step
all-reduces the histogram before chi-square (burst-split mode): by default inside
the evaluation kernel (its last block stores into every rank's inbox over NVLink,
--collective peer), or with ncclAllReduce + a finalize launch (--collective nccl).
--cfg 4 is the GA population: a step is one batch of 1024 parameter vectors over 1e5 bursts x 16
repeats in ONE launch (gsm_run_batch); at N > 1 the population is split across the ranks with no
collective (population mode) and the burst data is replicated (strong scaling of the batch).

  value     device-timed (CUDA events on the launching stream), inputs resident in HBM
  e2e       the same step through the public API with host buffers: set_nstates/set_params/
            setBurstBd/run_kernel + wait + read chi2 and histogram from pinned host memory
  roofline  HBM roofline of the evaluation kernel (algorithmic bytes 36 B/burst + 16 B/photon)
            plus the instruction-issue figure the north star asks for (`issue`)
  cpu_baseline  the checker (oracle/_ref = the reference's own kernel source built for the
            host, else the C port) timed on this box's host cores on a bounded sample

--impl reference times only that CPU path (rank 0), same metric/config.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from gsmfretda_b200 import synth  # noqa: E402

NB, LEFT_PAD = 200, 2
SEED = 20261104
# algorithmic thread-instructions per burst-sample, frozen from SURVEY.md §8(d) before optimisation
ALGO_OPS = {1: 400.0, 2: 610.0, 3: 1230.0, 4: 610.0, 5: 980.0}
try:   # roofline.json holds the frozen constants; the literals above are the same numbers
    with open(os.path.join(ROOT, "roofline.json")) as _f:
        ALGO_OPS = {int(k): float(v) for k, v in json.load(_f)["algorithmic_thread_ops_per_sample"].items()}
except Exception:   # noqa: BLE001
    pass


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


def ncu_traffic(workload: str, mode: str):
    """DRAM bytes per launch of the evaluation kernel from the committed ncu capture, or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            e = json.load(f).get(f"{workload}/{mode}")
        return (float(e["bytes"]), e["source"]) if e else (None, None)
    except Exception:
        return None, None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.proc = None
        self.lines = []   # (arrival time, csv line)
        self.t_begin = None
        self.t_end = None

    def mark_begin(self):
        self.t_begin = time.perf_counter()

    def mark_end(self):
        self.t_end = time.perf_counter()

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append((time.perf_counter(), ln.strip()))

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        inside = [ln for (t, ln) in self.lines
                  if self.t_begin is not None and self.t_begin - 0.05 <= t <= (self.t_end or t) + 0.15]
        window = "timed regions"
        if len(inside) < 3:   # a very short run: fall back to everything since the warm-up
            inside = [ln for (_, ln) in self.lines]
            window = "warm-up + timed regions (timed regions shorter than 3 samples)"
        for ln in inside:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": float(max(smax)) if smax else None,
                "samples": len(sm), "window": window, "reasons": sorted(reasons)}


class CpuPath:
    """The reference's CPU path for `cpu_baseline` and the `--impl reference` arm: the checker
    (oracle/_ref = the reference's own kernel source built for the host, else the C port) over a
    bounded sample of the workload, with all the host threads OpenMP gives it."""

    def __init__(self, w, sample_bursts, R):
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import oracle_py
        self.w = w
        self.O = oracle_py.Oracle()
        # every host core this process may run on, whatever OMP_NUM_THREADS says (torchrun sets it to 1)
        try:
            self.cores = len(os.sched_getaffinity(0))
        except AttributeError:
            self.cores = os.cpu_count() or self.O.num_threads()
        self.use_ref = oracle_py.have_ref()
        self.R = R
        self.note = ""
        if self.use_ref:
            self.Rf = oracle_py.Ref()
            self.R = 1 << (int(R).bit_length() - 1)  # mc::set_reSampleTimes rounds down to 2^k (mc.cu:316)
            if self.R != R:
                self.note = f"; the reference needs a power-of-two reSampleTimes: {R} -> {self.R}"
        self.sample_bursts = sample_bursts
        self.b = w.bursts(n_bursts=sample_bursts)
        self.ohist = self.O.observed_hist(self.b, NB)
        probe_n = max(1, min(sample_bursts, 2000))
        self.once(probe_n)
        self.rate = probe_n / max(self.once(probe_n), 1e-6)   # bursts per second, for sizing

    def once(self, nb_):
        w, O = self.w, self.O
        t0 = time.perf_counter()
        if self.use_ref:
            e = self.Rf.mc_kernel(self.b, w.n_states, w.args[0], self.R, SEED & 0xFFFFFFFF, 1, N=nb_, nthreads=self.cores)
            h = O.hist(e, NB)                     # streamWorker's host histogram + chi-square
            O.chi2(self.ohist, h, NB, self.R, LEFT_PAD, w.n_states)
        else:
            O.eval_chi2(self.b, w.n_states, w.args[0], self.R, SEED & 0xFFFFFFFF, 1, self.ohist, NB, LEFT_PAD, 0, nb_,
                        nthreads=self.cores)
        return time.perf_counter() - t0

    def timed(self, seconds_hint):
        """About `seconds_hint` seconds of work -> (burst-samples, seconds, evaluations, bursts each)."""
        nb_ = int(max(1, min(self.sample_bursts, self.rate * seconds_hint)))
        dt1 = self.once(nb_)
        reps = int(max(1, min(64, round(seconds_hint / max(dt1, 1e-6)))))  # whole evaluations, back to back
        dt = dt1 if reps == 1 else sum(self.once(nb_) for _ in range(reps))
        return nb_ * self.R * reps, dt, reps, nb_

    def describe(self, samples, dt, reps, nb_):
        src = ("reference mc_kernel source built for the host (oracle/_ref)" if self.use_ref
               else "C port (oracle/pda_oracle.c)")
        return {"unit": "burst-samples/s", "cores": self.cores, "kind": "reference" if self.use_ref else "port",
                "sample": f"{reps} evaluation(s) of {nb_} bursts x {self.R} repeats of {self.w.name} ({samples} "
                          f"burst-samples, {dt:.2f} s, {src} + host histogram + chi-square, OpenMP over samples{self.note})"}


def cpu_leg(w, sample_bursts, R, seconds_hint=12.0):
    """Time the reference's CPU path on a bounded sample -> dict for `cpu_baseline`."""
    cp = CpuPath(w, sample_bursts, R)
    samples, dt, reps, nb_ = cp.timed(seconds_hint)
    out = cp.describe(samples, dt, reps, nb_)
    out["value"] = samples / dt
    return out


def reference_on_gpu(cfg, bursts, timeout=90.0):
    """Second baseline of SURVEY.md 8(d): the reference's own `class mc` (unmodified mc.cu built
    for sm_100a, oracle/_ref/libref_gpu.so) through its own API on this GPU, in a process of its
    own (its destructor resets the device), at the reference's default reSampleTimes = 4."""
    exe = os.path.join(ROOT, "oracle", "refgpu_run.py")
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libref_gpu.so")):
        return {"unavailable": "oracle/_ref/libref_gpu.so not built"}
    try:
        r = subprocess.run([sys.executable, exe, "--cfg", str(cfg), "--bursts", str(bursts), "--repeats", "4",
                            "--evals", "1", "--time", "10"], capture_output=True, text=True, timeout=timeout)
        for ln in r.stdout.splitlines():
            if ln.startswith("{"):
                d = json.loads(ln)
                return {"value": d["burst_samples_per_s_with_host_hist"], "unit": "burst-samples/s",
                        "kernel_plus_d2h_only": d["burst_samples_per_s_kernel_d2h"],
                        "ms_per_evaluation": 1e3 * (d["kernel_plus_d2h_s_per_eval"] + d["host_hist_s_per_eval"]),
                        "sample": f"{d['evals_timed']} evaluations of {d['bursts']} bursts x {d['repeats']} repeats "
                                  f"(the reference's default reSampleTimes) through mc::run_kernel + D2H of N*R floats "
                                  f"+ host histogram (numpy bincount standing in for boost::histogram)"}
        return {"unavailable": f"no result line (rc {r.returncode})"}
    except subprocess.TimeoutExpired:
        return {"unavailable": f"the reference did not return within {timeout:.0f} s"}
    except Exception as e:   # noqa: BLE001
        return {"unavailable": repr(e)}


def run_reference(args, w, rank):
    if rank != 0:
        return
    # every step is a bounded sample of the workload, sized so that the whole run stays within
    # a few minutes whatever K is (about 120 s of CPU work in total)
    hint = min(5.0, max(0.05, 120.0 / max(1, args.steps + args.warmup)))
    cp = CpuPath(w, min(w.n_bursts, 100000), w.R)
    tot_samples, tot_s, last = 0, 0.0, None
    for i in range(args.warmup + args.steps):
        r = cp.timed(hint)
        if i >= args.warmup:
            tot_samples += r[0]; tot_s += r[1]; last = r
    val = tot_samples / tot_s
    cb = cp.describe(*last)
    cb["sample"] = f"per step: {cb['sample']}; {args.steps} steps, {tot_samples} burst-samples in {tot_s:.1f} s"
    cb["value"] = val
    out = {"impl": "reference", "metric": "burst-samples/s", "value": val, "unit": "burst-samples/s",
           "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": 1e3 * tot_s / max(1, args.steps), "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "f32 walk + f64 binomial", "data": "synthetic",
           "config": {"workload": w.name, "bursts": w.n_bursts, "repeats": w.R, "states": w.n_states,
                      "param_vectors": 1, "note": "CPU path of the reference on a bounded sample per step"},
           "cpu_baseline": cb,
           "e2e": {"value": val, "unit": "burst-samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cfg", type=int, default=2)
    ap.add_argument("--bursts", type=int, default=None, help="override bursts per GPU")
    ap.add_argument("--repeats", type=int, default=None)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--mode", default="replay", choices=["replay", "normal", "reference"],
                    help="replay: bit-exact draw contract (FP64 binomial); normal: FP32 fast arithmetic")
    ap.add_argument("--collective", default="peer", choices=["peer", "nccl"],
                    help="burst-split histogram reduction at N > 1: peer = fused into the evaluation kernel "
                         "(stores into every rank's inbox over NVLink), nccl = ncclAllReduce + a finalize launch")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    w = synth.workload(args.cfg, n_bursts=args.bursts, R=args.repeats)
    if args.impl == "reference":
        run_reference(args, w, rank)
        return

    # stdout carries exactly ONE line, the JSON: anything a library prints there meanwhile (NCCL's
    # version banner, for one) is sent to stderr by pointing fd 1 at fd 2 until the line is written
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist
    from gsmfretda_b200 import capi

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # stdout carries the JSON line only
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- data: every rank owns one shard of a world x B burst dataset ------------------
    population = args.cfg == 4
    # burst-split: every rank its own shard; population: the same bursts on every rank
    b = w.bursts(seed=w.seed if population else w.seed + 1000 * rank)
    B, R, n = b.n_bursts, w.R, w.n_states
    arg0 = w.args[0]
    P_job = w.args.shape[0] if population else 1
    if population:
        from gsmfretda_b200 import sharding
        p0, p1 = sharding.split_population(P_job, world)[rank]
        args_local = np.ascontiguousarray(w.args[p0:p1])
    P_local = (p1 - p0) if population else 1
    mc = capi.Mc(local_rank, 2)
    mc.set_gpuid()
    mc.set_reSampleTimes(R)
    mc.set_seed(SEED)
    mc.set_hist(NB, LEFT_PAD)
    MODES = {"replay": capi.GSM_MODE_REPLAY, "normal": capi.GSM_MODE_NORMAL, "reference": capi.GSM_MODE_REFERENCE}
    mc.set_mode(MODES[args.mode])
    mc.init_bursts(b, g_offset=0 if population else rank * B)
    if world > 1 and not population:
        if args.collective == "nccl":
            idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
            if rank == 0:
                idt.copy_(torch.frombuffer(bytearray(capi.comm_unique_id()), dtype=torch.uint8))
            dist.broadcast(idt, 0)
            mc.comm_init(bytes(idt.cpu().numpy().tobytes()), rank, world)
        else:
            mine = torch.frombuffer(bytearray(mc.peer_export(world, 1)), dtype=torch.uint8).cuda()
            allh = [torch.empty_like(mine) for _ in range(world)]
            dist.all_gather(allh, mine)
            mc.peer_connect(b"".join(bytes(h.cpu().numpy().tobytes()) for h in allh), rank, world)
        oh = torch.from_numpy(mc.observed_hist()).cuda()
        dist.all_reduce(oh)  # counts < 2^24: exact in f32
        mc.set_observed_hist(oh.cpu().numpy())
    samples_rank = B * R * P_local
    samples_job = B * R * P_job if population else samples_rank * world

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    def step_resident():
        flush.zero_()
        torch.cuda.synchronize()
        if population:
            mc.run_batch(B, 0, n, args_local)   # the vectors (612 B each) travel with the launch
        else:
            mc.run_resident(B, 0)
        mc.sync(0)
        return mc.elapsed_ms(0)

    def step_e2e():
        if population:
            mc.run_batch(B, 0, n, args_local)
            while not mc.streamQuery(0):
                pass
            return mc.chi2v(0), mc.hist(0, P_local - 1)
        mc.set_nstates(n, 0)
        mc.set_params(n, 0, arg0)
        Ns = mc.setBurstBd(0, -1, 0)
        mc.run_kernel(Ns, 0)
        while not mc.streamQuery(0):
            pass
        return mc.chi2(0), mc.hist(0, 0)

    # resident parameters for the device-timed leg
    chi2_first, hist_first = mc.evaluate(n, arg0, sid=0)
    assert int(hist_first.sum()) == (B * R if population else samples_job), (int(hist_first.sum()), samples_job)
    clocks = ClockSampler(local_rank)
    clocks.start()              # nvidia-smi needs ~0.5 s to come up: start it before the warm-up
    for _ in range(max(args.warmup, 3)):
        step_resident()
    l0 = mc.launch_count()
    barrier()
    clocks.mark_begin()
    t_wall0 = time.perf_counter()
    ms = [step_resident() for _ in range(args.steps)]
    barrier()
    t_wall = time.perf_counter() - t_wall0
    launches = mc.launch_count() - l0
    dev_ms = float(sum(ms))

    # end to end through the public API, host buffers in and out
    for _ in range(max(args.warmup, 3)):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        chi2_last, hist_last = step_e2e()
    barrier()
    e2e_s = time.perf_counter() - t0
    clocks.mark_end()
    clk = clocks.stop()
    assert int(hist_last.sum()) == (B * R if population else samples_job)
    if population:
        assert chi2_last.shape[0] == P_local and np.isfinite(chi2_last).all()
        chi2_last = float(chi2_last[0])

    # the same step in the other arithmetic (replay <-> normal), device-timed, for the record
    other = "normal" if args.mode == "replay" else "replay"
    mc.set_mode(MODES[other])
    k_other = max(3, min(args.steps, 50))
    for _ in range(3):
        step_resident()
    barrier()
    ms_other = float(sum(step_resident() for _ in range(k_other)))
    barrier()

    # measured issue peak (IMAD / FFMA chains on every SM), right after the timed regions
    peak_imad, peak_ffma = mc.issue_peak(0), mc.issue_peak(1)

    if world > 1:
        t = torch.tensor([dev_ms, e2e_s, t_wall, ms_other], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms, e2e_s, t_wall, ms_other = (float(x) for x in t.cpu())

    if rank == 0:
        peak, peak_src, sm_max = measured_peaks()
        ms_per_step = dev_ms / args.steps
        value = samples_job * args.steps / (dev_ms * 1e-3)
        algo_bytes = b.algorithmic_bytes() + (612 * P_local if population else 0)
        achieved = algo_bytes / (ms_per_step * 1e-3) / 1e9
        traffic, traffic_src = ncu_traffic(w.name, args.mode) if (args.bursts is None and args.repeats is None) else (None, None)
        f_clk = (clk.get("sm_mhz") or sm_max) * 1e6
        issue_peak = 148 * 128 * f_clk
        per_gpu = value / world
        out = {
            "metric": "burst-samples/s", "value": value, "unit": "burst-samples/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong" if population else "weak", "vs_baseline": None,
            "dtype": ("f32 walk + f64 binomial (replay arithmetic), u32/u64 histogram" if args.mode == "replay" else
                      "f32 walk + f64 binomial (replay arithmetic) on the reference's Sobol64/xor128 generators" if args.mode == "reference"
                      else "f32 (normal mode: hardware lg2/ex2/rcp, f32 binomial inversion), u32/u64 histogram"),
            "mode": args.mode,
            "data": "synthetic",
            "config": {"workload": w.name, "bursts_per_gpu": B, "photons_per_gpu": b.n_photons,
                       "repeats": R, "states": n, "param_vectors": P_job, "hist_bins": NB,
                       "sharding": "single GPU" if world == 1 else
                                   (f"population split x{world}, no collective" if population else
                                    (f"burst-split x{world}, histogram all-reduce fused into the evaluation kernel "
                                     f"(peer stores over NVLink, one launch per step)" if args.collective == "peer" else
                                     f"burst-split x{world}, NCCL all-reduce of the histogram + finalize launch")),
                       "l2": "256 MiB buffer written between timed iterations (L2 flush)",
                       "timing": "CUDA events on the launching stream per step, summed; max over ranks",
                       "photon_layout": "packed 4 B/photon" if mc.is_packed() else "raw 16 B/photon"},
            "param_evals_per_s": args.steps * P_job / (dev_ms * 1e-3),
            "other_mode": {"mode": other, "value": samples_job * k_other / (ms_other * 1e-3), "unit": "burst-samples/s",
                           "ms_per_step": ms_other / k_other, "steps": k_other},
            "wall_s_timed_region": t_wall,
            "e2e": {"value": samples_job * args.steps / e2e_s, "unit": "burst-samples/s",
                    "h2d_bytes_per_step": 612 * P_local, "d2h_bytes_per_step": 16 + P_local * ((NB + 2) * 8 + 4),
                    "param_evals_per_s": args.steps * P_job / e2e_s,
                    "ms_per_step": 1e3 * e2e_s / args.steps,
                    "transport": ("vectors by cudaMemcpyAsync from pinned memory, results by one DMA to pinned memory" if population else
                                  "the 612-byte vector rides in the launch parameters, the kernel's last block writes "
                                  "flags | histogram | chi-square into mapped pinned host memory: no copy on the stream"),
                    "api": ("Mc.run_batch/streamQuery + chi2v()/hist()" if population else
                            "Mc.set_nstates/set_params/setBurstBd/run_kernel/streamQuery + chi2()/hist()")},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": algo_bytes,
                         "note": "the kernel is instruction-issue bound, not HBM bound: see `issue`"},
            "issue": {"algorithmic_thread_ops_per_sample": ALGO_OPS[args.cfg],
                      "achieved_thread_ops_per_s": per_gpu * ALGO_OPS[args.cfg],
                      "peak_thread_ops_per_s": issue_peak, "frac": per_gpu * ALGO_OPS[args.cfg] / issue_peak,
                      "sm_clock_hz": f_clk, "note": "148 SMs x 128 lanes x SM clock during the run",
                      "measured_peak_imad_thread_ops_per_s": peak_imad, "measured_peak_ffma_thread_ops_per_s": peak_ffma,
                      "frac_of_measured_imad_peak": per_gpu * ALGO_OPS[args.cfg] / peak_imad if peak_imad > 0 else None},
            "clocks": clk,
            "chi2": chi2_last,
        }
        if not args.no_cpu and world == 1:   # the CPU baseline is reported at N = 1 only
            cb = cpu_leg(w, min(B, 100000), R, seconds_hint=12.0)
            out["cpu_baseline"] = cb
        else:
            out["cpu_baseline"] = None
        mc.close()
        mc = None
        if not args.no_cpu and world == 1:
            torch.cuda.synchronize()
            out["cpu_baseline"]["reference_on_gpu"] = reference_on_gpu(args.cfg, min(B, 100000))
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(out) + "\n").encode())
    if mc is not None:
        mc.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
