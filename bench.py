#!/usr/bin/env python
"""bench.py — throughput of the Monte-Carlo PDA evaluation (BASELINE.json metric:
burst-samples/s and param-evals/s) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--cfg 2] [--only main,cfg4,cfg5]

ONE JSON line.  The top level is the configuration BASELINE.json's metric is quoted on (config 2:
1e5 bursts x 100 repeats, one parameter vector; a "step" is one evaluation of the vector over the
whole burst set).  At N > 1 (torchrun) every rank holds its own 1e5-burst shard of one N x 1e5-burst
dataset (weak scaling) and each step reduces the histogram over all ranks before chi-square
(burst split): inside the evaluation kernel by default (its last block stores into
every rank's inbox over NVLink, --collective peer), or with ncclAllReduce + a finalize launch
(--collective nccl).  `configs` carries, from the same process, the other BASELINE configurations:

  cfg4  GA population: one step = 1024 parameter vectors over 1e5 bursts x 16 repeats in ONE launch
        per GPU (gsm_run_batch); at N > 1 the population is split across the ranks, no collective,
        the bursts replicated: STRONG scaling of param-evals/s
  cfg5  4-state model with background: ONE fixed 1e7-burst set (8 blocks of 1.25e6 bursts), split
        across the ranks by bursts, histogram reduced in the kernel: STRONG scaling; N = 1 holds all 1e7
  cfg3  3-state full K matrix, 1e6 bursts x 16 repeats (N = 1 only: BASELINE names one GPU)

For every leg:
  value     device-timed: CUDA events on the launching stream around each step's launch, the K steps
            queued back to back with a 256 MiB L2 flush between them (on the same stream, outside the
            events), inputs resident in HBM; max over ranks
  e2e       the same step through the public API with host buffers: set_nstates/set_params/
            setBurstBd/run_kernel + wait + read chi2 and histogram from pinned host memory
  roofline  the bound of this kernel is instruction issue: algorithmic thread-level INT32+FP32
            instructions (roofline.json, frozen before optimisation) per second against
            148 SMs x 128 lanes x the SM clock measured under load; the HBM figure rides in roofline.hbm
  cpu_baseline  the checker (oracle/_ref = the reference's own kernel source built for the
            host, else the C port) timed on this box's host cores on a bounded sample (N = 1 only)

At N > 1 the burst-split legs first prove themselves: rank 0 evaluates the concatenated dataset on
its one GPU and the histogram and chi-square BITS must equal the split evaluation's
(`parity_check: "bit-exact"`), else the run fails.  --impl reference times only the CPU path
(rank 0), same metric/config.
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

from gsmfretda_b200 import sharding, synth  # noqa: E402

NB, LEFT_PAD = 200, 2
# sizes of what crosses the bus per evaluation: one expanded parameter vector (struct ParamRaw) and the head of the
# result block (struct ResultHeader); gsmfret_c.cu holds static_asserts on both so these cannot drift again
PARAM_RAW_BYTES, RESULT_HEADER_BYTES = 616, 48
SEED = 20261104
CFG5_BLOCKS = 8          # the fixed 1e7-burst set of config 5 = 8 blocks of 1.25e6 bursts
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


def ncu_entry(workload: str, mode: str, bursts: int):
    """What the committed ncu capture of this workload's evaluation kernel says per launch
    (profiles/ncu_traffic.json): the entry for exactly this many bursts per GPU, else the one of the
    workload's nominal size when that is what runs."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            d = json.load(f)
        e = d.get(f"{workload}/{mode}@{int(bursts)}")
        if e is None and synth_nominal_bursts(workload) == int(bursts):
            e = d.get(f"{workload}/{mode}")
        return e
    except Exception:   # noqa: BLE001
        return None


def synth_nominal_bursts(workload: str):
    for c in (1, 2, 3, 4, 5):
        w = synth.workload(c)
        if w.name == workload:
            return w.n_bursts
    return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed regions (B200_PROFILING.md): one
    sampler for the process, read back per leg by time window."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.proc = None
        self.lines = []   # (arrival time, csv line)

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:   # noqa: BLE001
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append((time.perf_counter(), ln.strip()))

    def stop(self):
        if self.proc is None:
            return
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:   # noqa: BLE001
            self.proc.kill()

    def window(self, t_begin, t_end):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)    # let the sample that covers t_end arrive
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        inside = [ln for (t, ln) in list(self.lines) if t_begin - 0.05 <= t <= t_end + 0.15]
        window = "timed regions"
        if len(inside) < 3:   # a very short leg: widen to the second before it (its warm-up)
            inside = [ln for (t, ln) in list(self.lines) if t_begin - 1.0 <= t <= t_end + 0.15]
            window = "warm-up + timed regions (timed regions shorter than 3 samples)"
        sm, smax, pw, reasons = [], [], [], set()
        for ln in inside:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": float(max(smax)) if smax else None,
                "power_w": float(np.median(pw)) if pw else None,
                "samples": len(sm), "window": window, "reasons": sorted(reasons)}


# --------------------------------------------------------------------------------------------
# the reference's CPU path (cpu_baseline, --impl reference)
# --------------------------------------------------------------------------------------------
def pow2_passes(R):
    """mc::set_reSampleTimes keeps the highest power of two (/root/reference/src/mc.cu:316): R repeats
    on the reference are the passes of R's binary expansion (100 = 64 + 32 + 4)."""
    return [1 << k for k in range(int(R).bit_length() - 1, -1, -1) if (int(R) >> k) & 1]


class CpuPath:
    """The reference's CPU path for `cpu_baseline` and the `--impl reference` arm: the checker
    (oracle/_ref = the reference's own kernel source built for the host, else the C port) over a
    bounded sample of the workload, with all the host threads OpenMP gives it."""

    def __init__(self, w, b, R):
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
        self.R = int(R)
        self.passes = pow2_passes(R) if self.use_ref else [self.R]
        if self.use_ref:
            self.Rf = oracle_py.Ref()
        self.b = b
        self.sample_bursts = b.n_bursts
        self.ohist = self.O.observed_hist(self.b, NB)
        probe_n = max(1, min(self.sample_bursts, 2000))
        self.once(probe_n)
        self.rate = probe_n / max(self.once(probe_n), 1e-6)   # bursts per second, for sizing

    def once(self, nb_):
        w, O = self.w, self.O
        t0 = time.perf_counter()
        if self.use_ref:
            h = np.zeros(NB + 2, np.uint64)
            for i, r in enumerate(self.passes):   # all R repeats, in power-of-two passes
                e = self.Rf.mc_kernel(self.b, w.n_states, w.args[0], r, SEED & 0xFFFFFFFF, 1 + i, N=nb_, nthreads=self.cores)
                h = h + O.hist(e, NB)             # streamWorker's host histogram
            O.chi2(self.ohist, h, NB, self.R, LEFT_PAD, w.n_states)   # ... and chi-square
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
        how = (f" as passes of {'+'.join(str(p) for p in self.passes)} (the reference keeps a power of two per pass)"
               if self.use_ref and len(self.passes) > 1 else "")
        return {"unit": "burst-samples/s", "cores": self.cores, "kind": "reference" if self.use_ref else "port",
                "sample": f"{reps} evaluation(s) of {nb_} bursts x {self.R} repeats{how} of {self.w.name} ({samples} "
                          f"burst-samples, {dt:.2f} s, {src} + host histogram + chi-square, OpenMP over samples)"}


def cpu_leg(w, b, R, seconds_hint=12.0):
    """Time the reference's CPU path on a bounded sample -> dict for `cpu_baseline`."""
    cp = CpuPath(w, b, R)
    samples, dt, reps, nb_ = cp.timed(seconds_hint)
    out = cp.describe(samples, dt, reps, nb_)
    out["value"] = samples / dt
    return out


def reference_on_gpu(cfg, bursts, repeats, timeout=120.0):
    """Second baseline of SURVEY.md 8(d): the reference's own `class mc` (unmodified mc.cu built
    for sm_100a, oracle/_ref/libref_gpu.so) through its own API on this GPU, in a process of its
    own (its destructor resets the device): kernel + D2H of N*R floats + the host half of
    streamWorker (the reference's vendored boost::histogram + chi-square loop)."""
    exe = os.path.join(ROOT, "oracle", "refgpu_run.py")
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libref_gpu.so")):
        return {"unavailable": "oracle/_ref/libref_gpu.so not built"}
    try:
        r = subprocess.run([sys.executable, exe, "--cfg", str(cfg), "--bursts", str(bursts), "--repeats", str(repeats),
                            "--evals", "1", "--time", "5"], capture_output=True, text=True, timeout=timeout)
        for ln in r.stdout.splitlines():
            if ln.startswith("{"):
                d = json.loads(ln)
                return {"value": d["burst_samples_per_s_with_host_hist"], "unit": "burst-samples/s",
                        "kernel_plus_d2h_only": d["burst_samples_per_s_kernel_d2h"],
                        "ms_per_evaluation": 1e3 * (d["kernel_plus_d2h_s_per_eval"] + d["host_hist_s_per_eval"]),
                        "ms_host_half": 1e3 * d["host_hist_s_per_eval"], "stages_s": d.get("stages_s"),
                        "sample": f"{d['evals_timed']} evaluations of {d['bursts']} bursts x {d['repeats']} repeats "
                                  f"through mc::run_kernel + D2H of N*R floats + host histogram (the reference's vendored "
                                  f"boost::histogram) + chi-square"}
        return {"unavailable": f"no result line (rc {r.returncode}): {r.stderr.strip().splitlines()[-1:] or ''}"}
    except subprocess.TimeoutExpired as e:
        tail = (e.stderr.decode() if isinstance(e.stderr, bytes) else (e.stderr or "")).strip().splitlines()[-2:]
        return {"unavailable": f"the reference did not return within {timeout:.0f} s; last stages: {tail}"}
    except Exception as e:   # noqa: BLE001
        return {"unavailable": repr(e)}


def workload_config(w, B, nph, P_job, world, sharding_text, bursts_total):
    """The `config` object: the same function writes it for both arms, so their labels agree."""
    return {"workload": w.name, "bursts_per_gpu": int(B), "photons_per_gpu": int(nph), "bursts_total": int(bursts_total),
            "repeats": int(w.R),
            "states": int(w.n_states), "param_vectors": int(P_job), "hist_bins": NB, "n_gpus": int(world),
            "sharding": sharding_text,
            "l2": "GPU arm: 256 MiB buffer written between timed iterations (L2 flush); CPU arm: not applicable",
            "timing": "GPU arm: CUDA events on the launching stream around each step, summed, max over ranks; "
                      "CPU arm: wall clock around each bounded sample"}


def sharding_text(kind, world, collective):
    if world == 1:
        return "single GPU"
    if kind == "population":
        return f"population dealt over {world} ranks in shares of equal estimated cost, bursts replicated, no collective"
    red = ("histogram all-reduce fused into the evaluation kernel (peer stores over NVLink, one launch per step)"
           if collective == "peer" else "NCCL all-reduce of the histogram + finalize launch")
    return f"burst split x{world}, {red}"


def run_reference(args, w, rank):
    if rank != 0:
        return
    # every step is a bounded sample of the workload, sized so that the whole run stays within
    # a few minutes whatever K is (about 120 s of CPU work in total)
    hint = min(5.0, max(0.05, 120.0 / max(1, args.steps + args.warmup)))
    b = w.bursts(n_bursts=min(w.n_bursts, 100000))
    cp = CpuPath(w, b, w.R)
    tot_samples, tot_s, last = 0, 0.0, None
    for i in range(args.warmup + args.steps):
        r = cp.timed(hint)
        if i >= args.warmup:
            tot_samples += r[0]; tot_s += r[1]; last = r
    val = tot_samples / tot_s
    cb = cp.describe(*last)
    cb["sample"] = f"per step: {cb['sample']}; {args.steps} steps, {tot_samples} burst-samples in {tot_s:.1f} s"
    cb["value"] = val
    kind = "population" if args.cfg == 4 else "split"
    out = {"impl": "reference", "metric": "burst-samples/s", "value": val, "unit": "burst-samples/s",
           "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": 1e3 * tot_s / max(1, args.steps), "higher_is_better": True,
           "scaling": "strong" if kind == "population" else "weak", "vs_baseline": None,
           "dtype": "f32 walk + f64 binomial", "data": "synthetic",
           "config": workload_config(w, b.n_bursts, b.n_photons, w.args.shape[0] if kind == "population" else 1, args.gpus,
                                     sharding_text(kind, args.gpus, args.collective),
                                     b.n_bursts * (1 if kind == "population" else args.gpus)),
           "note": "the reference's CPU path on this box's host cores, every step a bounded sample of the workload "
                   "(all repeats, a share of the bursts); it does not use the GPUs, whatever --gpus says",
           "cpu_baseline": cb,
           "e2e": {"value": val, "unit": "burst-samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    print(json.dumps(out), flush=True)


# --------------------------------------------------------------------------------------------
# the GPU arm
# --------------------------------------------------------------------------------------------
class Env:
    """What every leg needs: the rank, torch, the process group, one clock sampler, one flush buffer."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        from gsmfretda_b200 import capi
        self.torch, self.dist, self.capi, self.args = torch, dist, capi, args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback")
        torch.cuda.set_device(self.local_rank)
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # stdout carries the JSON line only
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank))
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2
        self.clocks = ClockSampler(self.local_rank)
        self.clocks.start()          # nvidia-smi needs ~0.5 s to come up: long before the first timed region
        self.MODES = {"replay": capi.GSM_MODE_REPLAY, "normal": capi.GSM_MODE_NORMAL, "reference": capi.GSM_MODE_REFERENCE}

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, *vals):
        if self.world == 1:
            return [float(v) for v in vals]
        t = self.torch.tensor(list(vals), dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(x) for x in t.cpu()]

    def gather(self, vals):
        """-> [world, len(vals)] float64 on every rank."""
        v = self.torch.tensor(list(vals), dtype=self.torch.float64, device="cuda")
        if self.world == 1:
            return v.cpu().numpy()[None, :]
        out = [self.torch.empty_like(v) for _ in range(self.world)]
        self.dist.all_gather(out, v)
        return np.stack([o.cpu().numpy() for o in out])

    def make_mc(self, w, b, g_offset, split, mode, collective, max_P=1, streams=2):
        """A `class mc` with the leg's bursts resident; at N > 1 and split=True the ranks are joined for
        the histogram exchange and the observed histogram becomes the whole dataset's."""
        capi, torch, dist = self.capi, self.torch, self.dist
        mc = capi.Mc(self.local_rank, streams)
        mc.set_gpuid()
        mc.set_reSampleTimes(w.R)
        mc.set_seed(SEED)
        mc.set_hist(NB, LEFT_PAD)
        mc.set_mode(self.MODES[mode])
        t_init = time.perf_counter()
        mc.init_bursts(b, g_offset=g_offset)      # returns when the packed, indexed arrays are resident
        mc.init_ms = 1e3 * (time.perf_counter() - t_init)
        if split and self.world > 1:
            if collective == "nccl":
                idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
                if self.rank == 0:
                    idt.copy_(torch.frombuffer(bytearray(capi.comm_unique_id()), dtype=torch.uint8))
                dist.broadcast(idt, 0)
                mc.comm_init(bytes(idt.cpu().numpy().tobytes()), self.rank, self.world)
            else:
                mine = torch.frombuffer(bytearray(mc.peer_export(self.world, max_P)), dtype=torch.uint8).cuda()
                allh = [torch.empty_like(mine) for _ in range(self.world)]
                dist.all_gather(allh, mine)
                mc.peer_connect(b"".join(bytes(h.cpu().numpy().tobytes()) for h in allh), self.rank, self.world)
            oh = torch.from_numpy(mc.observed_hist()).cuda()
            dist.all_reduce(oh)  # counts < 2^24: exact in f32
            mc.set_observed_hist(oh.cpu().numpy())
        return mc


def bits(x):
    return int(np.float32(x).view(np.uint32))


def parity_check(env, w, mode, collective, split_result, make_whole):
    """The burst split proves itself on the real devices: rank 0 evaluates the whole (concatenated)
    dataset on its one GPU, global burst ids 0..; histogram and chi-square bits must equal what the
    split evaluation over the N ranks returned.  -> text for the JSON line, raises otherwise."""
    ok = 1
    detail = ""
    if env.rank == 0:
        whole = make_whole()
        mc1 = env.make_mc(w, whole, 0, False, mode, collective)
        c1, h1 = mc1.evaluate(w.n_states, w.args[0], sid=0)
        mc1.close()
        cs, hs = split_result
        if not (np.array_equal(h1, hs) and bits(c1) == bits(cs)):
            ok = 0
            detail = f"split chi2 {cs!r} vs one-GPU {c1!r}; {int((h1 != hs).sum())} histogram bins differ"
        else:
            detail = (f"histogram ({NB} + 2 bins, {int(h1.sum())} samples) and chi-square bits of the {env.world}-rank split equal "
                      f"rank 0's one-GPU evaluation of the concatenated {whole.n_bursts} bursts")
    ok = int(env.max_over_ranks(-ok)[0] == -1)
    if not ok:
        raise SystemExit(f"parity check failed: {detail}")
    return detail


def time_leg(env, cfg, w, b, kind, steps, warmup, mode, collective, g_offset=0, whole_fn=None, other_mode=True,
             result_on_ranks=None, sustained_s=0.0):
    """One configuration on this rank's GPU -> dict (complete on rank 0).
    kind 'split': the bursts are this rank's share of one dataset; 'population': the vectors are."""
    torch, capi = env.torch, env.capi
    world, rank = env.world, env.rank
    population = kind == "population"
    B, R, n = b.n_bursts, w.R, w.n_states
    arg0 = w.args[0]
    P_job = w.args.shape[0] if population else 1
    if population:   # shares of about equal cost: the vectors differ a lot in dwell segments per burst
        mine = sharding.deal_population([sharding.vector_cost(n, a) for a in w.args], world)[rank]
        args_local = np.ascontiguousarray(w.args[mine])
    P_local = len(mine) if population else 1
    mc = env.make_mc(w, b, g_offset, not population, mode, collective)
    tot_B = int(env.gather([B])[:, 0].sum()) if not population else B
    samples_job = B * R * P_job if population else tot_B * R

    def launch():
        if population:
            mc.run_batch(B, 0, n, args_local)   # the vectors (616 B each) travel with the launch
        else:
            mc.run_resident(B, 0)

    ext = torch.cuda.ExternalStream(mc.stream_ptr(0))

    def resident_steps(k):
        """k steps queued back to back on the library's stream: flush | event | launch | event.
        (A batch stages its vectors through one pinned buffer, so gsm_run_batch waits for the stream
        first: those steps are timed one by one with the library's own events around copy + kernel.)"""
        if population:
            ms_ = []
            for _ in range(k):
                with torch.cuda.stream(ext):
                    env.flush.zero_()
                launch()
                mc.sync(0)
                ms_.append(mc.elapsed_ms(0))
            return ms_
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(k)]
        with torch.cuda.stream(ext):
            for e0, e1 in ev:
                env.flush.zero_()
                e0.record()
                launch()
                e1.record()
        mc.sync(0)
        return [e0.elapsed_time(e1) for e0, e1 in ev]

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

    # resident parameters for the device-timed leg; the first evaluation also feeds the parity check
    chi2_first, hist_first = mc.evaluate(n, arg0, sid=0)
    assert int(hist_first.sum()) == (B * R if population else samples_job), (int(hist_first.sum()), samples_job)
    parity = None
    if world > 1 and not population and whole_fn is not None:
        parity = parity_check(env, w, mode, collective, (chi2_first, hist_first), whole_fn)

    resident_steps(max(warmup, 3))
    l0 = mc.launch_count()
    env.barrier()
    t_begin = time.perf_counter()
    ms = resident_steps(steps)
    env.barrier()
    t_wall = time.perf_counter() - t_begin
    launches = mc.launch_count() - l0
    dev_ms = float(sum(ms))

    # end to end through the public API, host buffers in and out
    for _ in range(max(warmup, 3)):
        step_e2e()
    env.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        chi2_last, hist_last = step_e2e()
    env.barrier()
    e2e_s = time.perf_counter() - t0
    t_end = time.perf_counter()
    clk = env.clocks.window(t_begin, t_end)
    assert int(hist_last.sum()) == (B * R if population else samples_job)
    if population:
        assert chi2_last.shape[0] == P_local and np.isfinite(chi2_last).all()
        chi2_last = float(chi2_last[0])

    # where the step's time goes at N > 1: %globaltimer stamps of the kernel's last block
    exchange = None
    if world > 1 and not population and collective == "peer":
        k2 = max(5, min(steps, 50))
        st = np.zeros((k2, 4), np.float64)
        for i in range(k2):
            with torch.cuda.stream(ext):
                env.flush.zero_()
            mc.run_resident(B, 0)
            mc.sync(0)
            t = mc.timing_ns(0).astype(np.int64)
            st[i] = (t - t[0]).astype(np.float64)
        allst = env.gather(st.reshape(-1)).reshape(world, k2, 4)
        to_seen = allst[:, :, 2] * 1e-3                       # last block start -> every rank's flag seen (us)
        waited = (allst[:, :, 2] - allst[:, :, 1]) * 1e-3     # own flags posted -> every rank's flag seen
        tail = (allst[:, :, 3] - allst[:, :, 2]) * 1e-3       # sum of the inboxes, chi-square, results out
        exchange = {"exchange_us": float(np.median(to_seen.min(axis=0) + tail.min(axis=0))),
                    "skew_us": float(np.median(waited.max(axis=0) - waited.min(axis=0))),
                    "post_us": float(np.median(allst[:, :, 1].min(axis=0) * 1e-3)),
                    "reduce_and_finalize_us": float(np.median(tail.min(axis=0))),
                    "steps": k2,
                    "how": "%globaltimer stamps written by the kernel's last block (start, flags posted, all ranks' flags "
                           "seen, done), host-synchronised steps; exchange_us = per step the smallest over ranks of "
                           "(start -> all seen) + (all seen -> done): the rank that arrives last waits for nobody, so that is "
                           "the cost of the exchange itself; skew_us = per step the spread over ranks of (posted -> all "
                           "seen): how long the first rank waited for the last; medians over the steps"}

    # the same step in the other arithmetic (replay <-> normal), device-timed, for the record
    other = None
    if other_mode:
        om = "normal" if mode == "replay" else "replay"
        mc.set_mode(env.MODES[om])
        k_other = max(3, min(steps, 50))
        resident_steps(3)
        env.barrier()
        ms_other = float(sum(resident_steps(k_other)))
        env.barrier()
        (ms_other,) = env.max_over_ranks(ms_other)
        other = {"mode": om, "value": samples_job * k_other / (ms_other * 1e-3), "unit": "burst-samples/s",
                 "ms_per_step": ms_other / k_other, "steps": k_other}
        mc.set_mode(env.MODES[mode])

    # the same resident launch back to back for >= sustained_s seconds: what a GA run (hours of launches) sees once
    # the clocks have settled, beside the short burst-clock headline above
    sustained = None
    if sustained_s > 0 and not population:
        (dev_ms_max,) = env.max_over_ranks(dev_ms)
        ms_ref = dev_ms_max / steps
        k_b = max(10, int(250.0 / ms_ref))                       # launches per event pair: about 0.25 s
        n_b = max(2, int(np.ceil(sustained_s * 1e3 / (k_b * ms_ref))))
        env.barrier()
        ts0 = time.perf_counter()
        ms_b = []
        for _ in range(n_b):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(ext):
                e0.record()
            for _ in range(k_b):
                mc.run_resident(B, 0)
            with torch.cuda.stream(ext):
                e1.record()
            mc.sync(0)
            ms_b.append(e0.elapsed_time(e1))
        env.barrier()
        ts1 = time.perf_counter()
        clk_s = env.clocks.window(ts0, ts1)
        tot_ms, last_ms = env.max_over_ranks(float(sum(ms_b)), float(np.mean(ms_b[-max(1, n_b // 4):])))
        v_s = samples_job * n_b * k_b / (tot_ms * 1e-3)
        sustained = {"value": v_s, "unit": "burst-samples/s", "ms_per_step": tot_ms / (n_b * k_b), "steps": n_b * k_b,
                     "seconds": tot_ms * 1e-3, "ms_per_step_last_quarter": last_ms / k_b,
                     "ratio_to_burst": v_s / (samples_job * steps / (dev_ms_max * 1e-3)),
                     "sm_mhz_median": clk_s.get("sm_mhz"), "power_w": clk_s.get("power_w"), "reasons": clk_s.get("reasons"),
                     "clock_samples": clk_s.get("samples"),
                     "how": f"{n_b} x {k_b} resident launches queued back to back on the library's stream, CUDA events around "
                            "every group, max over ranks; no L2 flush between launches (a GA run has none either), so unlike "
                            "the headline the photon words can stay in the 126 MB L2 where they fit"}

    peak_imad, peak_ffma = mc.issue_peak(0), mc.issue_peak(1)
    packed = mc.is_packed()
    init_ms = mc.init_ms
    mc.close()

    per_rank = env.gather([dev_ms / steps, clk.get("sm_mhz") or 0.0, clk.get("power_w") or 0.0, 1.0 if clk["reasons"] else 0.0])
    dev_ms, e2e_s, t_wall = env.max_over_ranks(dev_ms, e2e_s, t_wall)
    if rank != 0:
        return None

    peak, peak_src, sm_max = measured_peaks()
    ms_per_step = dev_ms / steps
    value = samples_job * steps / (dev_ms * 1e-3)
    algo_bytes = b.algorithmic_bytes() + (PARAM_RAW_BYTES * P_local if population else 0)
    hbm_achieved = algo_bytes / (float(per_rank[0, 0]) * 1e-3) / 1e9
    ent = ncu_entry(w.name, mode, B) if env.args.repeats is None else None
    f_clk = (clk.get("sm_mhz") or sm_max) * 1e6
    issue_peak = 148 * 128 * f_clk
    per_gpu = value / world
    ops = ALGO_OPS[cfg]
    out = {
        "metric": "burst-samples/s", "value": value, "unit": "burst-samples/s", "n_gpus": world,
        "steps": steps, "warmup": max(warmup, 3), "ms_per_step": ms_per_step,
        "higher_is_better": True,
        "scaling": "strong" if (population or cfg == 5) else "weak",
        "param_evals_per_s": steps * P_job / (dev_ms * 1e-3),
        "mode": mode,
        "dtype": ("f32 walk + f64 binomial (replay arithmetic), u32/u64 histogram" if mode == "replay" else
                  "f32 walk + f64 binomial (replay arithmetic) on the reference's Sobol64/xor128 generators" if mode == "reference"
                  else "f32 (normal mode: hardware lg2/ex2/rcp, f32 binomial inversion), u32/u64 histogram"),
        "config": workload_config(w, B, b.n_photons, P_job, world, sharding_text(kind, world, collective), tot_B),
        "photon_layout": "packed 4 B/photon" if packed else "raw 16 B/photon",
        "wall_s_timed_region": t_wall,
        "e2e": {"value": samples_job * steps / e2e_s, "unit": "burst-samples/s",
                "h2d_bytes_per_step": PARAM_RAW_BYTES * P_local,
                "d2h_bytes_per_step": RESULT_HEADER_BYTES + P_local * ((NB + 2) * 8 + 4),
                "param_evals_per_s": steps * P_job / e2e_s,
                "ms_per_step": 1e3 * e2e_s / steps,
                "transport": ("vectors by cudaMemcpyAsync from pinned memory, results by one DMA to pinned memory" if population else
                              f"the {PARAM_RAW_BYTES}-byte vector rides in the launch parameters, the kernel's last block writes "
                              "flags | histogram | chi-square into mapped pinned host memory: no copy on the stream"),
                "api": ("Mc.run_batch/streamQuery + chi2v()/hist()" if population else
                        "Mc.set_nstates/set_params/setBurstBd/run_kernel/streamQuery + chi2()/hist()")},
        "gpu_launches": int(launches),
        "roofline": {"bound": "issue", "achieved": per_gpu * ops, "peak": issue_peak,
                     "unit": "thread-instructions/s (INT32+FP32, algorithmic)", "frac": per_gpu * ops / issue_peak,
                     "algorithmic_thread_ops_per_sample": ops,
                     "peak_source": "148 SMs x 128 lanes x the median SM clock nvidia-smi saw during this leg",
                     "sm_clock_hz": f_clk,
                     "measured_peak_imad": peak_imad, "measured_peak_ffma": peak_ffma,
                     "frac_of_measured_imad_peak": per_gpu * ops / peak_imad if peak_imad > 0 else None,
                     "traffic": ent.get("bytes") if ent else None, "traffic_source": ent.get("source") if ent else None,
                     "executed": ({k: ent[k] for k in ("warp_instructions", "lanes_per_instruction", "issue_active_pct") if k in ent}
                                  if ent else None),
                     "hbm": {"bound": "hbm", "achieved": hbm_achieved, "peak": peak, "unit": "GB/s", "frac": hbm_achieved / peak,
                             "peak_source": peak_src, "algorithmic_bytes_per_launch": algo_bytes,
                             "note": "36 B/burst + 16 B/photon over the kernel's launch time: far from the HBM bound"}},
        "clocks": clk,
        "chi2": float(chi2_last),
    }
    out["init_ms"] = {"value": init_ms, "what": "gsm_init_data on this rank's bursts, host wall clock: H2D of the raw int64 "
                      "arrays from pageable memory, the pack and index kernels, the observed histogram (one-time, mc::init_data_gpu)",
                      "bursts": B, "photons": int(b.n_photons)}
    if other:
        out["other_mode"] = other
    if sustained:
        out["sustained"] = sustained
    if world > 1:
        out["per_rank"] = {"ms_per_step": [float(x) for x in per_rank[:, 0]], "sm_mhz": [float(x) for x in per_rank[:, 1]],
                           "power_w": [float(x) for x in per_rank[:, 2]], "any_throttle_reason": [bool(x) for x in per_rank[:, 3]]}
    if exchange:
        out.update(exchange)
    if parity:
        out["parity_check"] = "bit-exact"
        out["parity_detail"] = parity
    return out


def gen_blocks(w, ids, bursts_each):
    """Blocks `ids` of a blocked dataset (block j: seed w.seed + 1000 j), generated on threads (numpy
    releases the GIL in the passes that matter)."""
    from concurrent.futures import ThreadPoolExecutor
    ids = list(ids)
    if len(ids) <= 1:
        return [w.bursts(n_bursts=bursts_each, seed=w.seed + 1000 * j) for j in ids]
    with ThreadPoolExecutor(min(4, len(ids))) as ex:
        return list(ex.map(lambda j: w.bursts(n_bursts=bursts_each, seed=w.seed + 1000 * j), ids))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cfg", type=int, default=2, help="the top-level configuration (BASELINE quotes its metric on 2)")
    ap.add_argument("--only", default=None, help="comma list of legs: main,cfg3,cfg4,cfg5 (default: all when --cfg is 2 "
                                                 "and no size override, else main)")
    ap.add_argument("--bursts", type=int, default=None, help="override bursts per GPU (cfg 5: bursts of the whole set)")
    ap.add_argument("--repeats", type=int, default=None)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline legs")
    ap.add_argument("--sustained", type=float, default=5.0,
                    help="seconds of back-to-back resident launches after the headline of the main and cfg5 legs (0 = off)")
    ap.add_argument("--mode", default="replay", choices=["replay", "normal", "reference"],
                    help="replay: bit-exact draw contract (FP64 binomial); normal: FP32 fast arithmetic")
    ap.add_argument("--collective", default="peer", choices=["peer", "nccl"],
                    help="burst-split histogram reduction at N > 1: peer = fused into the evaluation kernel "
                         "(stores into every rank's inbox over NVLink), nccl = ncclAllReduce + a finalize launch")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    w = synth.workload(args.cfg, n_bursts=args.bursts, R=args.repeats)
    if args.impl == "reference":
        run_reference(args, w, rank)
        return
    if args.only:
        legs = [x.strip() for x in args.only.split(",") if x.strip()]
    elif args.cfg == 2 and args.bursts is None and args.repeats is None:
        legs = ["main", "cfg4", "cfg5", "cfg3"]
    else:
        legs = ["main"]

    # stdout carries exactly ONE line, the JSON: anything a library prints there meanwhile (NCCL's
    # version banner, for one) is sent to stderr by pointing fd 1 at fd 2 until the line is written
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    env = Env(args)
    world = env.world
    cpu = (not args.no_cpu) and world == 1      # the CPU baselines are reported at N = 1 only
    t_start = time.perf_counter()

    def say(msg):
        if rank == 0:
            print(f"[bench +{time.perf_counter() - t_start:6.1f} s] {msg}", file=sys.stderr, flush=True)

    def blocked_leg(cfg, wl, steps, whole_small=None):
        """Config 5's way: ONE fixed set of CFG5_BLOCKS blocks, rank r holds a contiguous range of blocks."""
        each = wl.n_bursts // CFG5_BLOCKS
        if world > CFG5_BLOCKS:
            return ({"skipped": f"the fixed set has {CFG5_BLOCKS} blocks; {world} ranks asked"} if rank == 0 else None), None
        j0, j1 = sharding.split_population(CFG5_BLOCKS, world)[env.rank]
        b = synth.concat(gen_blocks(wl, range(j0, j1), each))
        say(f"{wl.name}: blocks [{j0}, {j1}) of {each} bursts resident on rank 0")
        parity = None
        if world > 1:   # bit-exactness of the split on the real devices, on a reduced copy of the set (same seeds)
            small = max(1, min(each, 20000))
            ws = synth.workload(cfg, n_bursts=small * CFG5_BLOCKS, R=wl.R)
            bs = synth.concat(gen_blocks(ws, range(j0, j1), small))
            mcs = env.make_mc(ws, bs, j0 * small, True, args.mode, args.collective)
            res = mcs.evaluate(ws.n_states, ws.args[0], sid=0)
            parity = parity_check(env, ws, args.mode, args.collective, res,
                                  lambda: synth.concat(gen_blocks(ws, range(CFG5_BLOCKS), small)))
            mcs.close()
        d = time_leg(env, cfg, wl, b, "split", steps, args.warmup, args.mode, args.collective, g_offset=j0 * each,
                     other_mode=False, sustained_s=args.sustained)
        if d is not None:
            d["config"]["dataset"] = (f"one fixed set of {CFG5_BLOCKS} blocks x {each} bursts (block j: seed {wl.seed} + 1000 j), "
                                      f"the same {CFG5_BLOCKS * each} bursts at every N; rank r holds a contiguous range of blocks")
            if parity:
                d["parity_check"] = "bit-exact"
                d["parity_detail"] = parity + " (the same blocks cut to 20000 bursts each)"
        return d, b

    out = None
    extra = {}
    cpu_jobs = []     # (where to put it, workload, bursts, R, seconds)
    for leg in legs:
        if leg == "main":
            kind = "population" if args.cfg == 4 else "split"
            if args.cfg == 5:
                out, b = blocked_leg(5, w, args.steps)
            else:
                # burst split: every rank its own shard of one world x B set; population: the same bursts everywhere
                b = w.bursts(seed=w.seed if kind == "population" else w.seed + 1000 * env.rank)
                B = b.n_bursts
                whole = (lambda: synth.concat([w.bursts(seed=w.seed + 1000 * r) for r in range(world)]))
                out = time_leg(env, args.cfg, w, b, kind, args.steps, args.warmup, args.mode, args.collective,
                               g_offset=0 if kind == "population" else env.rank * B, whole_fn=whole,
                               sustained_s=args.sustained)
            say(f"main leg ({w.name}) done")
            if cpu and rank == 0:
                nb_cpu = min(b.n_bursts, 100000)
                cpu_jobs.append((out, w, b if b.n_bursts == nb_cpu else b.slice(0, nb_cpu), w.R, 12.0))
        elif leg in ("cfg3", "cfg4", "cfg5"):
            c = int(leg[3])
            if c == args.cfg and "main" in legs:
                continue
            wl = synth.workload(c)
            k = max(3, min(args.steps, {3: 100, 4: 20, 5: 50}[c]))
            if c == 3:
                if world > 1:
                    continue          # BASELINE names one GPU for config 3
                bl = wl.bursts()
                d = time_leg(env, 3, wl, bl, "split", k, args.warmup, args.mode, args.collective, other_mode=False)
            elif c == 4:
                bl = wl.bursts()
                d = time_leg(env, 4, wl, bl, "population", k, args.warmup, args.mode, args.collective, other_mode=False)
            else:
                d, bl = blocked_leg(5, wl, k)
            say(f"{leg} leg ({wl.name}) done")
            if rank == 0:
                extra[leg] = d
                if cpu and isinstance(d, dict) and "value" in d and bl is not None:
                    nb_cpu = min(bl.n_bursts, 100000)
                    cpu_jobs.append((d, wl, bl.slice(0, nb_cpu), wl.R, 4.0))
            del bl
        else:
            raise SystemExit(f"unknown leg {leg!r}")

    env.clocks.stop()
    if rank == 0:
        if out is None:     # --only without main: promote the first leg measured
            first = next(iter(extra))
            out = extra.pop(first)
        out["vs_baseline"] = None       # BASELINE.md holds no published number for this metric
        out["data"] = "synthetic"
        if extra:
            out["configs"] = extra
        for (dst, wl, bl, R, secs) in cpu_jobs:
            dst["cpu_baseline"] = cpu_leg(wl, bl, R, seconds_hint=secs)
        out.setdefault("cpu_baseline", None)
        if cpu and "main" in legs:
            env.torch.cuda.synchronize()
            nb_ref = min(w.n_bursts, 100000)
            rg = {}
            for r_ in (4, 32, 128):   # the reference's default reSampleTimes and two larger ones
                rg[f"R{r_}"] = reference_on_gpu(args.cfg, nb_ref if r_ == 4 else min(nb_ref, 20000), r_)
            rg["R64"] = {"unavailable": "the reference never returns at reSampleTimes = 64: its kernel maps a thread to its burst "
                                        "with tidx >> (int)log2((double)reSampleTimes) on the device (mc.cu:101), CUDA's log2(64.0) "
                                        "is one ulp below 6, the shift becomes 5 and half the threads walk bursts that do not exist "
                                        "(tools/log2_probe.cu, profiles/r02_refgpu_R64_hang.txt)"}
            out["cpu_baseline"]["reference_on_gpu"] = rg["R4"]
            out["cpu_baseline"]["reference_on_gpu_by_repeats"] = rg
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(out) + "\n").encode())
    if world > 1:
        env.dist.destroy_process_group()


if __name__ == "__main__":
    main()
