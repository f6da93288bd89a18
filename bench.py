#!/usr/bin/env python
"""Benchmark of the direct-sum (exhaustive) step -- BASELINE.json's metric:
pairwise interactions/s (and steps/s) at 1M bodies, FP32, on 1/2/4/8 B200.

    python bench.py --gpus N --steps K --warmup W            (N > 1: under torchrun)
    python bench.py --impl reference ...                     (the reference's CPU code)

One "step" = one fused force + leapfrog launch over all N^2 ordered pairs
(incl. j = i, as the reference counts them).  Prints ONE JSON line on rank 0.

Headline keys (value, e2e, roofline, cpu_baseline, clocks, config) describe BASELINE
config 3 (1M-body Plummer sphere, FP32).  After the headline timed region the same
process measures the other BASELINE configs and reports them under "configs":
    fp32_1M_spread   general-mass FP32 kernel (no equal-mass specialisation)
    fp32_1M_uniform  uniform-cube workload, FP32
    fp64_100k        config 2: 100k bodies, FP64 direct sum
    bh_1M            config 4: 1M-body Barnes-Hut, accuracy against the direct sum
    bh_1M_uniform    the same on the uniform cube (tree cost depends on the distribution)
    bh_16M           config 5: 16M-body Barnes-Hut (world == 8 only, or --bh16m)
    small_n          N = 128 .. 100k: us/step and GB/s on the 64 / 128 B/body of a step
plus "parity" (the bench workload against the CPU oracle, the ranks against each other
and against one GPU; non-zero exit when it fails), "ref_gpu_baseline" (the reference's own
CUDA programs on this GPU) and the CPU baselines of BASELINE.md section 4.
"""
import argparse
import hashlib
import json
import os
import re
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)

FLOP_PER_INTERACTION = 20.0   # BASELINE.md section 3 convention
SM_COUNT, FP32_LANES, SM_MAX_MHZ = 148, 128, 1965.0
FP32_PEAK_TFLOPS = SM_COUNT * FP32_LANES * 2 * SM_MAX_MHZ * 1e6 / 1e12   # 74.4 nominal
FP64_PEAK_TFLOPS = FP32_PEAK_TFLOPS / 2                                  # 37.2 nominal
# tools/microbench.cu on this pool's B200 (profiles/r1_microbench_pipe_rates.jsonl)
FFMA_LANES_MEASURED = 123.09      # FFMA lane-ops / clk / SM
DFMA_LANES_MEASURED = 62.2        # DFMA lane-ops / clk / SM
FP64_PEAK_MEASURED_TFLOPS = SM_COUNT * DFMA_LANES_MEASURED * 2 * SM_MAX_MHZ * 1e6 / 1e12   # 36.2
MIX_CLK_F32 = 12.5                # 6 FFMA2 + 1 MUFU: clk per interaction-warp per sub-partition
BH_FP64_OPS_PER_TEST = 17         # FP64 warp instructions per child record the walk tests


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        return 6650.0, "fallback of the profiling recipe (MEASURED_PEAKS.json absent)"


# ------------------------------------------------------------- workloads ----
def plummer(n, seed=17):
    """Aarseth-Henon-Wielen Plummer sphere: M = 1, G = 1, a = 1, r <= 10a,
    COM position/velocity removed.  Returns pos_xyzm[n,4], vel[n,3] (float64)."""
    rng = np.random.default_rng(seed)
    r = np.empty(0)
    while r.size < n:
        u = rng.uniform(1e-12, 1.0, n)
        c = 1.0 / np.sqrt(u ** (-2.0 / 3.0) - 1.0)
        r = np.concatenate([r, c[c <= 10.0]])
    r = r[:n]

    def iso(m):
        z = rng.uniform(-1, 1, m)
        ph = rng.uniform(0, 2 * np.pi, m)
        s = np.sqrt(1 - z * z)
        return np.stack([s * np.cos(ph), s * np.sin(ph), z], axis=1)

    pos = r[:, None] * iso(n)
    q = np.empty(0)
    while q.size < n:
        x = rng.uniform(0, 1, n)
        y = rng.uniform(0, 0.1, n)
        q = np.concatenate([q, x[y < x * x * (1 - x * x) ** 3.5]])
    q = q[:n]
    speed = q * np.sqrt(2.0) * (1 + r * r) ** -0.25
    vel = speed[:, None] * iso(n)
    pos -= pos.mean(axis=0)
    vel -= vel.mean(axis=0)
    pm = np.concatenate([pos, np.full((n, 1), 1.0 / n)], axis=1)
    return np.ascontiguousarray(pm), np.ascontiguousarray(vel)


def uniform_cube(n, seed=17):
    """SURVEY 8d "Uniform": positions U(-1,1)^3, velocities 0, m = 1/N, seed 17."""
    rng = np.random.default_rng(seed)
    pm = np.concatenate([rng.uniform(-1.0, 1.0, (n, 3)), np.full((n, 1), 1.0 / n)], axis=1)
    return np.ascontiguousarray(pm), np.zeros((n, 3))


def fixture_like(n, seed=17):
    """The recipe of the reference's tests/*_g1.csv fixtures: numpy seed 17,
    pos = randn, vel = randn minus the mass-weighted mean, mass 0.2.  tests/100k_bodies.csv is
    absent from the reference checkout (.MISSING_LARGE_BLOBS); this re-synthesises it."""
    rs = np.random.RandomState(seed)
    pos = rs.randn(n, 3)
    vel = rs.randn(n, 3)
    vel -= vel.mean(axis=0)
    pm = np.concatenate([pos, np.full((n, 1), 0.2)], axis=1)
    return np.ascontiguousarray(pm), np.ascontiguousarray(vel)


def make_workload(distribution, n):
    return uniform_cube(n) if distribution == "uniform" else plummer(n)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons (B200_PROFILING.md), one process for the whole
    bench; every timed region reads the samples that fell inside its wall-clock window."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        self.index, self.proc, self.samples = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
            t0 = time.time()
            while not self.samples and time.time() - t0 < 5.0:   # first sample takes ~1 s
                time.sleep(0.05)
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.samples.append((time.time(), ln.strip()))

    def stop(self):
        if self.proc:
            self.proc.terminate()
            self.t.join(timeout=2)

    def window(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        inside = [ln for (t, ln) in self.samples if t0 <= t <= t1 + 0.15]
        if not inside:   # region shorter than the sampling period: nearest sample
            inside = [min(self.samples, key=lambda s: abs(s[0] - t1))[1]] if self.samples else []
        mhz, mx, pw, reasons = [], [], [], set()
        for ln in inside:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                mhz.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, val in zip(self.NAMES, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(mhz)) if mhz else None,
                "sm_max_mhz": max(mx) if mx else None,
                "power_w": float(np.median(pw)) if pw else None,
                "samples": len(mhz), "reasons": sorted(reasons)}


# ------------------------------------------------------------- CPU legs ----
def cpu_reference_rate(pm, sample_n, evals, threads, warmup=1):
    """Times the reference's own acceleration() (openmp/exhaustive.c:166-200, compiled in
    place into oracle/_ref with -O3) on the first sample_n bodies of the workload; falls
    back to the oracle port when oracle/_ref is absent.  Returns (interactions/s, kind, sec)."""
    import oracle as O
    sub = np.ascontiguousarray(pm[:sample_n])
    acc = np.zeros((sample_n, 3))
    ref_path = os.path.join(O.REFDIR, "libref_exh_omp_fast.so")
    if os.path.exists(ref_path):
        lib = O.load_ref("exh_omp_fast")
        lib.ref_set_threads(threads)
        fn, kind = (lambda: lib.ref_acceleration(sample_n, sub, acc)), "reference"
    else:
        lib = O.load_oracle(fast=True)
        fn, kind = (lambda: lib.orc_acceleration_mt(sample_n, sub, acc, threads)), "port"
    for _ in range(max(1, warmup)):
        fn()  # warm-up (page faults, thread pool)
    t0 = time.perf_counter()
    for _ in range(evals):
        fn()
    sec = time.perf_counter() - t0
    return evals * float(sample_n) ** 2 / sec, kind, sec


def write_csv(path, pm, vel):
    """the reference's input format (pos, vel, mass per line, %.17g) through the C host library"""
    from conftest import load_package
    h = load_package().load_host_library()
    h.nbh_write_bodies(path.encode(), len(pm), np.ascontiguousarray(pm), np.ascontiguousarray(vel))


def run_program(cmd, timeout):
    """Runs one of the reference's stand-alone programs in its own process group (killed as a
    group on timeout: nothing is left behind holding the GPU); returns its own 'Elapsed wall
    time' (seconds) or None, plus the process wall time."""
    import signal
    t0 = time.perf_counter()
    p = subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                         start_new_session=True)
    try:
        out, err = p.communicate(timeout=timeout)
    except subprocess.TimeoutExpired:
        try:
            os.killpg(p.pid, signal.SIGKILL)
        except OSError:
            pass
        p.communicate()
        return None, time.perf_counter() - t0, "timeout after %d s (killed)" % timeout
    wall = time.perf_counter() - t0
    m = re.search(r"Elapsed wall time: ([0-9.]+)", out)
    if not m:
        return None, wall, (err or out)[-200:]
    return float(m.group(1)), wall, None


def cpu_program_baselines(pm, vel, sample_n, threads):
    """the reference's exh_mp as shipped (Makefile flags = -O0) and
    -O3, run as programs on a CSV of the first sample_n bodies for ONE step (2 evaluations),
    timed by their own 'Elapsed wall time' line."""
    import oracle as O
    out = {}
    with tempfile.TemporaryDirectory() as td:
        inp = os.path.join(td, "in.csv")
        write_csv(inp, pm[:sample_n], vel[:sample_n])
        for key, exe in (("exh_mp_as_shipped_O0", "exh_mp_shipped"), ("exh_mp_O3", "exh_mp_fast")):
            path = os.path.join(O.REFDIR, exe)
            if not os.path.exists(path):
                out[key] = {"unavailable": "oracle/_ref/%s not built" % exe}
                continue
            sec, wall, err = run_program([path, inp, "0", "0.01", "0.01", os.path.join(td, "o.csv"),
                                          str(threads)], 300)
            if sec is None:
                out[key] = {"unavailable": err}
                continue
            out[key] = {"value": 2.0 * sample_n * sample_n / sec, "unit": "interactions/s",
                        "cores": threads, "kind": "reference", "elapsed_wall_time_s": sec,
                        "sample": "openmp/exhaustive.c program on the first %d bodies, 1 step = 2 "
                                  "evaluations, its own 'Elapsed wall time'" % sample_n}
    return out


def bh_cpu_baseline(pm, vel, sample_n, threads, shipped=False):
    """The reference's barnes-hut-omp (openmp/barnes-hut.c, compiled in place into oracle/_ref)
    on the first sample_n bodies, one step = two force evaluations; timed by its own 'Elapsed
    wall time' line (the CSV read is outside it, as in the reference)."""
    import oracle as O
    exe = os.path.join(O.REFDIR, "barnes_hut_omp_shipped" if shipped else "barnes_hut_omp_fast")
    if not os.path.exists(exe):
        return None
    with tempfile.TemporaryDirectory() as td:
        inp = os.path.join(td, "in.csv")
        write_csv(inp, pm[:sample_n], vel[:sample_n])
        sec, wall, err = run_program([exe, inp, "0", "0.01", "0.01", os.path.join(td, "o.csv"),
                                      str(threads)], 600)
    if sec is None:
        return {"unavailable": err}
    return {"value": 2.0 * sample_n / sec, "unit": "body force evaluations/s", "cores": threads,
            "kind": "reference", "elapsed_wall_time_s": sec,
            "sample": "barnes-hut-omp (openmp/barnes-hut.c, %s) on %d bodies, 1 step "
                      "(2 evaluations), %.2f s by its own clock"
                      % ("as shipped, -O0" if shipped else "-O3", sample_n, sec)}


def ref_gpu_baseline(pm100k, vel100k, pm1m, vel1m, ours_fp64_rate, ours_bh_ms):
    """BASELINE.md section 4, last bullet: the reference's OWN CUDA programs (cuda/exhaustive.cu,
    cuda/barnes-hut.cu, nvcc -arch=sm_100a, unmodified) on this B200, timed by their own
    'Elapsed wall time'.  exhaustive-cuda at 100k bodies with grid 98 x block 1024, for 2 and
    for 6 steps: the difference removes the CUDA start-up its clock includes."""
    import oracle as O
    out = {}
    exh = os.path.join(O.REFDIR, "exhaustive_cuda_ref")
    bhx = os.path.join(O.REFDIR, "barnes_hut_cuda_ref")
    with tempfile.TemporaryDirectory() as td:
        if os.path.exists(exh):
            n = len(pm100k)
            inp = os.path.join(td, "in100k.csv")
            write_csv(inp, pm100k, vel100k)
            t = {}
            for steps, end in ((2, "0.02"), (6, "0.06")):
                sec, wall, err = run_program([exh, inp, "0", end, "0.01", os.path.join(td, "o.csv"),
                                              "98", "1024"], 300)
                t[steps] = sec
                if sec is None:
                    out["exhaustive_cuda"] = {"unavailable": err}
                    break
            if t.get(2) and t.get(6):
                per_eval = (t[6] - t[2]) / 4.0
                rate = float(n) * n / per_eval if per_eval > 0 else None
                out["exhaustive_cuda"] = {
                    "bodies": n, "grid": 98, "block": 1024, "elapsed_2_steps_s": t[2],
                    "elapsed_6_steps_s": t[6], "s_per_evaluation": per_eval,
                    "value": rate, "unit": "interactions/s",
                    "value_incl_startup": 3.0 * n * n / t[2],
                    "ours_fp64_same_n": ours_fp64_rate,
                    "ours_over_reference": (ours_fp64_rate / rate) if (rate and ours_fp64_rate) else None,
                    "what": "cuda/exhaustive.cu:273-338 (FP64, pow(r2,-1.5), 4 launches + 4 syncs per "
                            "step) compiled nvcc -arch=sm_100a, its own 'Elapsed wall time'"}
        else:
            out["exhaustive_cuda"] = {"unavailable": "oracle/_ref/exhaustive_cuda_ref not built"}
        if os.path.exists(bhx) and pm1m is not None:
            n = len(pm1m)
            inp = os.path.join(td, "in1m.csv")
            write_csv(inp, pm1m, vel1m)
            sec, wall, err = run_program([bhx, inp, "0", "0.01", "0.01", os.path.join(td, "o.csv")], 240)
            if sec is None:
                out["barnes_hut_cuda"] = {"unavailable": err}
            else:
                out["barnes_hut_cuda"] = {
                    "bodies": n, "elapsed_1_step_s": sec, "s_per_evaluation": sec / 2.0,
                    "value": 2.0 * n / sec, "unit": "body force evaluations/s",
                    "ours_ms_per_evaluation": ours_bh_ms,
                    "ours_over_reference": (sec / 2.0 * 1e3 / ours_bh_ms) if ours_bh_ms else None,
                    "what": "cuda/barnes-hut.cu:814-1186 compiled nvcc -arch=sm_100a, 1 step = 2 "
                            "evaluations + its per-step D2H and fprintf of every body, own clock"}
        elif pm1m is not None:
            out["barnes_hut_cuda"] = {"unavailable": "oracle/_ref/barnes_hut_cuda_ref not built"}
        else:
            out["barnes_hut_cuda"] = {
                "unavailable": "not run by default: cuda/barnes-hut.cu (nvcc -arch=sm_100a, unmodified) does not "
                               "terminate on a B200 -- 10k bodies x 1 step was killed after 240 s in round 2 "
                               "(gpurun_out/r2_refgpu.json; its lock-based insertion, add_ent :227-295, spins "
                               "on child slots inside one warp).  --ref-gpu-bh tries again under a timeout",
                "ours_ms_per_evaluation": ours_bh_ms}
    return out


def workload_config(args):
    """identical for both arms and every N (the driver compares it): nothing rank- or
    arm-dependent in here"""
    gen = "plummer seed 17" if args.distribution == "plummer" else "uniform cube seed 17"
    shape = "Plummer sphere" if args.distribution == "plummer" else "uniform cube U(-1,1)^3"
    return {"workload": "synthetic %d-body %s, exhaustive direct sum, %s, eps2=0.01, dt=0.01f"
                        % (args.bodies, shape, args.precision),
            "bodies": args.bodies, "precision": args.precision, "generator": gen,
            "l2": "flushed between timed steps (256 MiB write); positions are rewritten every step"}


def run_reference_arm(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores:
    acceleration() of openmp/exhaustive.c (-O3) on a bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    pm, _ = make_workload(args.distribution, args.bodies)
    sample_n = min(args.cpu_sample, args.bodies)
    value, kind, sec = cpu_reference_rate(pm, sample_n, args.steps, threads, warmup=args.warmup)
    full = float(args.bodies) ** 2 / value
    sample = ("acceleration() of openmp/exhaustive.c (-O3, FP64) on the first %d bodies of the "
              "workload, %d passes; rate extrapolated to %d bodies by N^2 (BASELINE.md section 4)"
              % (sample_n, args.steps, args.bodies))
    line = {
        "impl": "reference", "metric": "pairwise interactions/s", "value": value,
        "unit": "interactions/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": sec / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args),
        "sample_bodies": sample_n, "extrapolated": True,
        "step_is": "one acceleration() pass over the %d-body sample (%.3g interactions)"
                   % (sample_n, float(sample_n) ** 2),
        "s_per_full_step_extrapolated": full,
        "cpu_baseline": {"value": value, "unit": "interactions/s", "cores": threads, "kind": kind,
                         "sample": sample},
        "e2e": {"value": value, "unit": "interactions/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# -------------------------------------------------------------- GPU side ----
class Ctx:
    """rank plumbing shared by every measurement of one bench process"""

    def __init__(self, args):
        import torch
        from conftest import load_package
        self.torch, self.nb, self.args = torch, load_package(), args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        self.sampler = ClockSampler(self.local)
        if self.rank == 0:
            self.sampler.start()

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, vals, op="max"):
        t = self.torch.tensor([float(v) for v in vals], dtype=self.torch.float64, device="cuda")
        if self.dist is not None:
            ops = {"max": self.dist.ReduceOp.MAX, "min": self.dist.ReduceOp.MIN,
                   "sum": self.dist.ReduceOp.SUM}
            self.dist.all_reduce(t, op=ops[op])
        return [float(x) for x in t]

    def engine(self, pm, vel, **kw):
        """a world-rank engine; rank 0 creates the communicator id, torch.distributed carries it"""
        comm = self.nb.share_comm_id(self.dist, self.rank, device="cuda") if self.world > 1 else None
        return self.nb.Engine(pm, vel, device=self.local, rank=self.rank, world=self.world,
                              comm_id_bytes=comm, **kw)

    def ranks_agree(self, *arrays):
        """True iff every rank holds bit-identical arrays (hash, then min == max over ranks)"""
        h = hashlib.sha256()
        for a in arrays:
            h.update(np.ascontiguousarray(a).tobytes())
        v = int.from_bytes(h.digest()[:6], "little")   # 48 bits: exact in a double
        lo = self.reduce([v], "min")[0]
        hi = self.reduce([v], "max")[0]
        return lo == hi

    def timed(self, eng, steps, warmup):
        """W untimed + K timed single-step launches, L2 flushed before each, device time from
        the engine's CUDA events, max over ranks.  Returns (step ms total, force ms total,
        launches, clocks during the region)."""
        torch = self.torch
        for _ in range(warmup):
            eng.step(1)
        eng.sync()
        l0 = eng.stats().kernel_launches
        self.barrier()
        t0 = time.time()
        dev_ms = force_ms = 0.0
        for _ in range(steps):
            self.flush.zero_()                 # L2 flush: 256 MiB > 126 MB L2
            torch.cuda.synchronize()
            eng.step(1)
            st = eng.stats()                   # syncs the engine stream, reads its CUDA events
            dev_ms += st.last_step_ms
            force_ms += st.last_force_ms
        self.barrier()
        t1 = time.time()
        launches = eng.stats().kernel_launches - l0
        dev_ms, force_ms = self.reduce([dev_ms, force_ms], "max")
        clocks = self.sampler.window(t0, t1) if self.rank == 0 else None
        self.last_wall = t1 - t0
        return dev_ms, force_ms, int(launches), clocks

    def steps_for(self, eng, target_s, lo, hi):
        """how many steps fill ~target_s (same answer on every rank)"""
        eng.step(1)
        ms = max(self.reduce([eng.stats().last_step_ms], "max")[0], 1e-3)
        return int(min(hi, max(lo, round(target_s * 1e3 / ms))))

    def close(self):
        if self.rank == 0:
            self.sampler.stop()
        if self.dist is not None:
            self.dist.barrier()
            self.dist.destroy_process_group()


def direct_roofline(nb, prec, uniform_mass, my_rows, n_padded, force_ms_per_step, clocks):
    peak = FP32_PEAK_TFLOPS if prec == nb.PREC_FP32 else FP64_PEAK_TFLOPS
    per_launch_flops = FLOP_PER_INTERACTION * float(my_rows) * float(n_padded)
    ach = per_launch_flops / (force_ms_per_step * 1e-3) / 1e12
    roof = {"bound": "fma" if prec == nb.PREC_FP32 else "fp64", "achieved": ach, "peak": peak,
            "unit": "TFLOP/s", "frac": ach / peak, "traffic": None,
            "peak_source": ("nominal: 148 SM x 128 FP32 lanes x 2 x 1965 MHz" if prec == nb.PREC_FP32 else
                            "nominal: 148 SM x 64 FP64 lanes x 2 x 1965 MHz") +
                           " (MEASURED_PEAKS.json has no FMA-pipe figure; sm_max_mhz from it)",
            "flop_per_interaction": FLOP_PER_INTERACTION}
    if prec == nb.PREC_FP32:
        roof["peak_measured_ffma"] = SM_COUNT * FFMA_LANES_MEASURED * 2 * SM_MAX_MHZ * 1e6 / 1e12
        roof["ceiling_of_instruction_mix"] = (SM_COUNT * 4 * 32 / (MIX_CLK_F32 * (11.0 / 12.0 if uniform_mass else 1.0))
                                              * SM_MAX_MHZ * 1e6 * FLOP_PER_INTERACTION / 1e12)
    else:
        roof["peak_measured_dfma"] = FP64_PEAK_MEASURED_TFLOPS
        roof["frac_of_measured_dfma"] = ach / FP64_PEAK_MEASURED_TFLOPS
        # 16 (equal masses: 15) FP64-pipe operations + 1 MUFU per interaction at the measured 62.2
        # DFMA lanes/clk/SM
        roof["ceiling_of_instruction_mix"] = (SM_COUNT * DFMA_LANES_MEASURED / (15.0 if uniform_mass else 16.0)
                                              * SM_MAX_MHZ * 1e6 * FLOP_PER_INTERACTION / 1e12)
    roof["frac_of_instruction_mix_ceiling"] = ach / roof["ceiling_of_instruction_mix"]
    if clocks and clocks.get("sm_mhz"):
        roof["frac_at_observed_clock"] = ach / (peak * clocks["sm_mhz"] / SM_MAX_MHZ)
    bytes_per_launch = (64.0 if prec == nb.PREC_FP32 else 128.0) * my_rows
    hp, hs = hbm_peak()
    hbm_ach = bytes_per_launch / (force_ms_per_step * 1e-3) / 1e9
    roof["hbm_algorithmic_bytes_per_launch"] = bytes_per_launch
    roof["hbm"] = {"achieved": hbm_ach, "peak": hp, "unit": "GB/s", "frac": hbm_ach / hp, "peak_source": hs}
    return roof


def direct_record(ctx, what, pm, vel, prec, steps, warmup, tune_flags=0, target_s=None):
    """one direct-sum config, sharded over all ranks like the headline"""
    nb = ctx.nb
    n = len(pm)
    with ctx.engine(pm, vel, precision=prec, tune_flags=tune_flags) as eng:
        st0 = eng.stats()
        my_rows = st0.slice_end - st0.slice_begin
        if target_s:
            steps = ctx.steps_for(eng, target_s, 5, 200)
        dev_ms, force_ms, launches, clocks = ctx.timed(eng, steps, warmup)
        uni = eng.stats().uniform_mass
        variant = eng.stats().kernel_variant
    if ctx.rank != 0:
        return None
    return {"workload": what, "bodies": n, "precision": "fp32" if prec == nb.PREC_FP32 else "fp64",
            "n_gpus": ctx.world, "steps": steps, "warmup": warmup, "ms_per_step": dev_ms / steps,
            "value": steps * float(n) * n / (dev_ms * 1e-3), "unit": "interactions/s",
            "steps_per_s": steps / (dev_ms * 1e-3), "gpu_launches": launches, "clocks": clocks,
            "kernel": kernel_name(nb, prec, uni, variant),
            "roofline": direct_roofline(nb, prec, uni, my_rows, st0.n_padded, force_ms / steps, clocks)}


def kernel_name(nb, prec, uniform_mass, variant=None):
    if prec != nb.PREC_FP32:
        return ("direct_f64_kernel (variant %s%s): %d FP64-pipe ops + 1 MUFU per interaction"
                % (variant, ", equal masses: m_j hoisted" if uniform_mass else "", 15 if uniform_mass else 16))
    # nb_direct_f32.cu kVariants: 0..2 pack two j-bodies per register pair (direct_f32_kernel),
    # 3.. pack two i-bodies and take j as a broadcast scalar operand (direct_f32_ip_kernel)
    name = "direct_f32_kernel, j-packed f32x2" if variant is not None and variant < 3 else \
           "direct_f32_ip_kernel, i-packed f32x2 (8 i-bodies per thread, j broadcast from LDS.128)"
    return name + (" (variant %s), equal masses: m_j hoisted, 11 FMA-pipe ops + 1 MUFU per interaction" % variant
                   if uniform_mass else
                   " (variant %s), general masses: 12 FMA-pipe ops + 1 MUFU per interaction" % variant)


def e2e_loop(ctx, eng, pm, vel, reps):
    """nb_upload(_slice) + nb_step(1) + nb_gather(_slice) per step from pinned host buffers,
    wall clock, max over ranks.  With world > 1 every rank moves only its rows over PCIe."""
    torch = ctx.torch
    b, c = eng.io_slice()
    hp = torch.from_numpy(pm[b:b + c].copy()).pin_memory()
    hv = torch.from_numpy(vel[b:b + c].copy()).pin_memory()
    op = torch.empty_like(hp).pin_memory()
    ov = torch.empty_like(hv).pin_memory()

    def once():
        eng.upload_slice_raw(hp.data_ptr(), hv.data_ptr())
        eng.step(1)
        eng.gather_slice_raw(op.data_ptr(), ov.data_ptr())

    once()
    ctx.barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        once()
    ctx.barrier()
    wall = ctx.reduce([time.perf_counter() - t0], "max")[0]
    checksum = ctx.reduce([float(op[:, :3].abs().sum())], "sum")[0]
    nbytes_in = ctx.reduce([hp.numel() * 8 + hv.numel() * 8], "sum")[0]
    nbytes_out = ctx.reduce([op.numel() * 8 + ov.numel() * 8], "sum")[0]
    return wall, checksum, int(nbytes_in), int(nbytes_out)


def parity_rows(ctx, eng, pm, vel, prec, rows=32):
    """(i) of the parity record: accelerations of `rows` random bodies of the BENCH workload from
    the engine (all ranks, through nb_accelerations) against the CPU oracle's rows."""
    nb = ctx.nb
    n = len(pm)
    eng.upload(pm, vel)
    acc = eng.accelerations()
    if ctx.rank != 0:
        return None
    import oracle as O
    lib = O.load_oracle()
    src = pm.astype(np.float32).astype(np.float64) if prec == nb.PREC_FP32 else pm
    src = np.ascontiguousarray(src)
    idx = np.sort(np.random.default_rng(99).choice(n, size=min(rows, n), replace=False))
    worst = 0.0
    out = np.zeros((1, 3))
    for i in idx:
        lib.orc_acceleration_slice(n, src, int(i), int(i) + 1, out)
        den = np.linalg.norm(out[0])
        worst = max(worst, float(np.linalg.norm(acc[i] - out[0]) / (den if den > 0 else 1.0)))
    return worst


def parity_vs_single(ctx, pm, vel, steps, **kw):
    """(ii): `steps` leapfrog steps on all ranks vs the same on one GPU (rank 0 alone).
    Returns (ranks bitwise equal, max relative difference to the 1-rank state)."""
    from conftest import rel_err
    with ctx.engine(pm, vel, **kw) as e:
        e.step(steps)
        p, v = e.gather()
    same = ctx.ranks_agree(p, v)
    diff = None
    if ctx.rank == 0:
        with ctx.nb.Engine(pm, vel, device=ctx.local, **kw) as e1:
            e1.step(steps)
            p1, v1 = e1.gather()
        diff = float(max(rel_err(p, p1).max(), rel_err(v, v1).max()))
    ctx.barrier()
    return same, diff


def small_n_record(ctx):
    """The HBM / launch-latency regime: us per step and GB/s on the algorithmic
    64 B (FP32) / 128 B (FP64) per body of one step, K steps enqueued back to back."""
    nb = ctx.nb
    hp, hs = hbm_peak()
    rows = []
    for n, k in ((128, 400), (1000, 400), (10_000, 200), (100_000, 20)):
        pm, vel = fixture_like(n)
        for prec, nm, bpb in ((nb.PREC_FP32, "fp32", 64.0), (nb.PREC_FP64, "fp64", 128.0)):
            with nb.Engine(pm, vel, precision=prec, device=ctx.local) as e:
                e.step(5)
                e.sync()
                best = None
                for _ in range(3):
                    e.step(k)
                    ms = e.stats().last_step_ms / k
                    best = ms if best is None else min(best, ms)
            rows.append({"bodies": n, "precision": nm, "us_per_step": best * 1e3,
                         "interactions_per_s": float(n) * n / (best * 1e-3),
                         "hbm_gbs": bpb * n / (best * 1e-3) / 1e9,
                         "hbm_frac": bpb * n / (best * 1e-3) / 1e9 / hp})
    floor = min(r["us_per_step"] for r in rows if r["bodies"] == 128)
    return {"rows": rows, "launch_latency_floor_us": floor,
            "what": "nb_step(K) back to back on one GPU, CUDA events / K, best of 3; bytes = 64 B "
                    "(FP32) or 128 B (FP64) per body per step (read x,v; write x,v); the 128-body "
                    "row is the launch-latency floor of one fused step",
            "hbm_peak_gbs": hp, "hbm_peak_source": hs}


# ------------------------------------------------------- Barnes-Hut legs ----
def bh_record(ctx, what, pm, vel, accuracy=True, cpu_baseline=False, single_gpu_check=True):
    """BASELINE configs 4/5: Barnes-Hut (THETA = 0.5, FP64).  value = body force evaluations per
    second; roofline = FP64-pipe work of the walk (17 FP64 warp instructions per child record a
    warp tests) against the measured DFMA rate, plus the stages' algorithmic bytes per second
    against the measured HBM copy bandwidth."""
    from conftest import rel_err
    nb = ctx.nb
    n = len(pm)
    kw = dict(precision=nb.PREC_FP64, algorithm=nb.ALGO_BARNES_HUT)
    with ctx.engine(pm, vel, **kw) as eng:
        eng.step(2)
        steps = ctx.steps_for(eng, 1.5, 5, 200)
        dev_ms, _f, launches, clocks = ctx.timed(eng, steps, 3)
        # the same steps as ONE nb_step(steps) call: no host sync and no L2 flush between steps
        # (a step's working set -- positions, keys, the tree -- is larger than L2 from 1M bodies
        # up); with world > 1 the own-slice sort of step k+1 runs under the position all-gather
        # of step k, which the step-at-a-time loop above cannot show
        eng.upload(pm, vel)                    # the same stretch of the system's evolution
        eng.step(2 + 1 + 3)
        eng.sync()
        ctx.barrier()
        eng.step(steps)
        b2b_ms = ctx.reduce([eng.stats().last_step_ms], "max")[0] / steps
        wall, _chk, nin, nout = e2e_loop(ctx, eng, pm, vel, 3)
        p, v = eng.gather()
        same = ctx.ranks_agree(p, v) if ctx.world > 1 else None
        acc_stats = None
        if accuracy and n <= 2_000_000:
            eng.upload(pm, vel)
            a_bh = eng.accelerations()
    if accuracy and n <= 2_000_000:
        with ctx.engine(pm, vel, precision=nb.PREC_FP64) as d:
            a_dir = d.accelerations()
        err = rel_err(a_bh, a_dir)
        acc_stats = {"median": float(np.median(err)), "p99": float(np.percentile(err, 99)),
                     "max": float(err.max()),
                     "envelope_ok": bool(np.median(err) <= 4e-3 and np.percentile(err, 99) <= 2e-2
                                         and err.max() <= 1.2e-1),
                     "what": "per-body relative acceleration error vs the FP64 direct-sum kernel; "
                             "the reference's own THETA envelope is median 4e-3 / p99 2e-2 / max 1.2e-1"}
    # stage times: an engine that launches the ~75 kernels plainly (no graph) records stage events
    with ctx.engine(pm, vel, tune_flags=nb.TUNE_NO_GRAPH, **kw) as se:
        se.step(3)
        se.sync()
        ctx.barrier()
        se.step(1)
        sst = se.stats()
        plain = ctx.reduce([sst.last_force_ms, sst.last_step_ms], "max")
        stage_ms = list(se.tree_stats().stage_ms)
    trav_min = ctx.reduce([stage_ms[6]], "min")[0]
    stage_ms = ctx.reduce(stage_ms, "max")
    # tree statistics from a counting engine (the counting walk is ~10 % slower: never timed)
    with ctx.engine(pm, vel, tune_flags=nb.TUNE_COUNT_VISITS, **kw) as ce:
        ce.step(1)
        ti = ce.tree_stats()
    inter, visits, tests = ctx.reduce([ti.interactions, ti.node_visits, ti.child_tests], "sum")
    tests_max = ctx.reduce([ti.child_tests], "max")[0]
    vs1 = None
    if ctx.world > 1 and single_gpu_check:
        m = min(n, 200_000)
        same2, vs1 = parity_vs_single(ctx, pm[:m].copy(), vel[:m].copy(), 3, **kw)
        same = bool(same and same2)
    if ctx.rank != 0:
        return None
    hp, hs = hbm_peak()
    nodes = ti.nodes
    own = n / ctx.world
    names = ["keys", "sort", "merge", "order", "build", "com", "traversal"]
    # algorithmic bytes per stage per rank: keys read 32 + write 12 per body; sort
    # 8 passes x 32 (world > 1: keys and sort of the OWN slice only, then ceil(log2 W) merge passes
    # of 24 B over all keys); order 68; build 432 per cell (48 + the 384-byte child block);
    # centre of mass 360 per cell; traversal = the cells' child blocks once (384 per cell) + 132
    # per own body (body, index, v in/out, x out)
    sorted_here = own if ctx.world > 1 else n
    merge_passes = int(np.ceil(np.log2(ctx.world))) if ctx.world > 1 else 0
    stage_bytes = [44.0 * sorted_here, 256.0 * sorted_here, 24.0 * n * merge_passes, 68.0 * n,
                   432.0 * nodes, 360.0 * nodes, 384.0 * nodes + 132.0 * own]
    exchange_ms = stage_ms[7] if ctx.world > 1 else None
    stage_ms = stage_ms[:7]
    stages = {}
    for nm, ms, by in zip(names, stage_ms, stage_bytes):
        gbs = by / (ms * 1e-3) / 1e9 if ms > 0 else None
        stages[nm] = {"ms": ms, "algorithmic_bytes": by, "gbs": gbs, "hbm_frac": (gbs / hp) if gbs else None}
    if exchange_ms is not None: # not a kernel of ours: ncclAllGather of the positions (32 B/body)
        stages["exchange"] = {"ms": exchange_ms, "algorithmic_bytes": 32.0 * n,
                              "what": "position all-gather on the side stream, max over ranks, the "
                                      "wait for the slowest rank's traversal included"}
    trav_ms = stage_ms[6]
    stages["traversal"]["ms_min_over_ranks"] = trav_min
    # the same step of the plainly launched engine as a whole: launch_force (the wait for the
    # other ranks' cube maxima and sorted runs included) and nb_step (the position exchange too)
    stages["whole_evaluation_plain_launches"] = {"ms": plain[0]}
    stages["whole_step_plain_launches"] = {"ms": plain[1]}
    fp64_ach = BH_FP64_OPS_PER_TEST * 32.0 * tests_max * 2.0 / (trav_ms * 1e-3) / 1e12 if trav_ms > 0 else None
    total_stage = sum(stage_ms)
    total_bytes = sum(stage_bytes)
    rec = {
        "workload": what, "bodies": n, "precision": "fp64", "theta": 0.5, "n_gpus": ctx.world,
        "steps": steps, "ms_per_step": dev_ms / steps, "value": n * steps / (dev_ms * 1e-3),
        "unit": "body force evaluations/s", "gpu_launches": launches, "clocks": clocks,
        "timing": "one nb_step(1) per timed iteration, L2 flushed and the host synchronised in between",
        "back_to_back": {"ms_per_step": b2b_ms, "value": n / (b2b_ms * 1e-3),
                         "what": "the same steps as one nb_step(%d) call (no flush: the step's "
                                 "working set of %.0f MB exceeds the 126 MB L2)"
                                 % (steps, (ti.nodes * 464.0 + n * 120.0) / 1e6)},
        "tree": {"nodes": nodes, "depth": ti.depth, "interactions_per_body": inter / n,
                 "visits_per_body": visits / n, "child_tests_per_warp": tests / (n / 32.0)},
        "accuracy_vs_direct": acc_stats,
        "e2e": {"value": 3 * 2.0 * n / wall, "unit": "body force evaluations/s",
                "h2d_bytes_per_step": nin, "d2h_bytes_per_step": nout,
                "what": "nb_upload_slice + nb_step(1) + nb_gather_slice per step = 2 evaluations"},
        "roofline": {
            "bound": "fp64", "achieved": fp64_ach, "peak": FP64_PEAK_MEASURED_TFLOPS, "unit": "TFLOP/s",
            "frac": (fp64_ach / FP64_PEAK_MEASURED_TFLOPS) if fp64_ach else None,
            # traverse_kernel, 1M Plummer on one GPU: dram read 373.9 MB + write 54.2 MB per launch
            # (profiles/r2_bh_traverse_ncu_full_before.csv); other shapes were not captured
            "traffic": 428099072 if (n == 1_000_000 and ctx.world == 1 and "Plummer" in what) else None,
            "what": "traversal kernel (%.0f %% of the evaluation): %d FP64 warp instructions per child "
                    "record tested x 32 lanes x 2 flop, over the traversal's device time, against the "
                    "measured DFMA rate 62.2 lanes/clk/SM x 148 SM x 1965 MHz; a record costs the same "
                    "whether 1 or 32 lanes still need it" % (100.0 * trav_ms / max(total_stage, 1e-9),
                                                             BH_FP64_OPS_PER_TEST),
            "stages": stages,
            "hbm": {"achieved": total_bytes / (total_stage * 1e-3) / 1e9 if total_stage > 0 else None,
                    "peak": hp, "unit": "GB/s", "peak_source": hs,
                    "frac": (total_bytes / (total_stage * 1e-3) / 1e9 / hp) if total_stage > 0 else None,
                    "what": "algorithmic bytes of all stages / their summed device time (engine with "
                            "plain launches, stage events)"}},
        "parity": {"ranks_bitwise": same, "vs_single_gpu_200k_max_rel": vs1,
                   "what": "gathered state hashed on every rank (min == max over ranks); 3 steps at 200k "
                           "bodies on all ranks vs on one GPU (a body's interaction list does not depend "
                           "on who walks it: expected 0.0)"},
    }
    if cpu_baseline:
        threads = os.cpu_count() or 1
        rec["cpu_baseline"] = bh_cpu_baseline(pm, vel, min(n, 1_000_000), threads)
        rec["cpu_baseline_as_shipped"] = bh_cpu_baseline(pm, vel, min(n, 200_000), threads, shipped=True)
    return rec


def bh_parity_ok(rec):
    if rec is None:
        return True
    par = rec.get("parity") or {}
    ok = par.get("ranks_bitwise") in (None, True)
    if par.get("vs_single_gpu_200k_max_rel") is not None:
        ok = ok and par["vs_single_gpu_200k_max_rel"] == 0.0
    acc = rec.get("accuracy_vs_direct")
    if acc is not None:
        ok = ok and acc["envelope_ok"]
    return ok


# -------------------------------------------------------------- GPU arm ----
def run_gpu_arm(args):
    ctx = Ctx(args)
    torch, nb = ctx.torch, ctx.nb
    rank, world = ctx.rank, ctx.world
    prec = nb.PREC_FP32 if args.precision == "fp32" else nb.PREC_FP64
    n = args.bodies
    pm, vel = make_workload(args.distribution, n)
    if args.masses == "spread":
        pm[:, 3] = np.random.default_rng(18).uniform(0.5, 1.5, n) / n
    eng = ctx.engine(pm, vel, precision=prec, ctas_per_sm=args.ctas_per_sm, tune_flags=args.tune_flags)
    st0 = eng.stats()
    my_rows = st0.slice_end - st0.slice_begin

    # ---- kernel-resident throughput: inputs already in HBM -------------------------------
    dev_ms, force_ms, launches, clocks = ctx.timed(eng, args.steps, args.warmup)
    wall = ctx.last_wall
    value = float(args.steps) * float(n) * float(n) / (dev_ms * 1e-3)

    # ---- the same K steps enqueued back to back (one nb_step call, no L2 flush, no host sync
    # in between): what a simulation run sees; with world > 1 the all-gather of step k overlaps
    # the own-slice launch of step k+1 here.  Reported beside `value`, not instead of it.
    ctx.barrier()
    eng.step(args.steps)
    stb = eng.stats()
    b2b_ms = ctx.reduce([stb.last_step_ms], "max")[0] / args.steps

    # ---- end to end through the C ABI with host buffers -----------------------------------
    # one e2e step = propagation() of one time step from host arrays: H2D of positions and
    # velocities (pinned), kick+drift launch, closing force launch, D2H of the new state;
    # it evaluates 2 N^2 interactions, as the reference does for n_steps = 1.  With world > 1
    # every rank moves its own rows over PCIe (nb_upload_slice / nb_gather_slice), the rest
    # crosses NVLink.
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    e2e_wall, checksum, nin, nout = e2e_loop(ctx, eng, pm, vel, e2e_steps)
    e2e_value = e2e_steps * 2.0 * float(n) * float(n) / e2e_wall

    # ---- parity on the bench workload, outside every timed region --------------------------
    uni = eng.stats().uniform_mass
    rows_tol = 1e-5 if prec == nb.PREC_FP32 else 1e-12
    rows_err = parity_rows(ctx, eng, pm, vel, prec)
    eng.step(2)
    gp, gv = eng.gather()
    ranks_bitwise = ctx.ranks_agree(gp, gv) if world > 1 else None
    eng.close()
    vs_single, vs_tol = None, (2e-6 if prec == nb.PREC_FP32 else 1e-13) * 1e3
    if world > 1:
        m = min(n, 200_000)
        same2, vs_single = parity_vs_single(ctx, pm[:m].copy(), vel[:m].copy(), 3, precision=prec)
        ranks_bitwise = bool(ranks_bitwise and same2)
    parity = {"rows": 32, "max_rel": rows_err, "rows_tolerance": rows_tol,
              "ranks_bitwise": ranks_bitwise, "vs_single_gpu": vs_single,
              "vs_single_gpu_tolerance": vs_tol if world > 1 else None,
              "what": "nb_accelerations of 32 random bodies of the %d-body bench workload vs "
                      "orc_acceleration_slice (CPU oracle, FP64, float-rounded inputs for fp32); "
                      "ranks_bitwise: gathered state hashed on every rank; vs_single_gpu: 3 steps of "
                      "the first 200k bodies on all ranks vs on one GPU, max relative difference "
                      "(summation order differs: own slice first)" % n}
    ok = True
    if rank == 0:
        ok = rows_err is not None and rows_err <= rows_tol
        if world > 1:
            ok = ok and ranks_bitwise and vs_single is not None and vs_single <= vs_tol
    parity["ok"] = bool(ok)

    # ---- the other BASELINE configs -----------------------------------------------------------
    configs = {}
    only = set(args.only.split(",")) if args.only else None

    def want(name):
        return not args.no_configs and (only is None or name in only)

    if not args.no_configs:
        sub_steps = max(3, min(args.steps, 5))
        if want("fp32_1M_spread"):
            pm_s = pm.copy()
            pm_s[:, 3] = np.random.default_rng(18).uniform(0.5, 1.5, n) / n
            configs["fp32_1M_spread"] = direct_record(
                ctx, "synthetic %d-body Plummer sphere, masses U(0.5,1.5)/N: the general FP32 kernel" % n,
                pm_s, vel, nb.PREC_FP32, sub_steps, 3)
            del pm_s
        pu, vu = uniform_cube(n)
        if want("fp32_1M_uniform"):
            configs["fp32_1M_uniform"] = direct_record(
                ctx, "synthetic %d-body uniform cube U(-1,1)^3, v = 0, m = 1/N, FP32" % n,
                pu, vu, nb.PREC_FP32, sub_steps, 3)
        if want("fp64_100k"):
            p100, v100 = fixture_like(100_000)
            configs["fp64_100k"] = direct_record(
                ctx, "BASELINE config 2: 100k bodies (the recipe of the reference's fixtures: numpy seed 17, "
                     "randn, mass 0.2; tests/100k_bodies.csv itself is missing from the reference checkout), "
                     "FP64 direct sum", p100, v100, nb.PREC_FP64, 20, 3, target_s=1.5)
        nbh = min(n, 1_000_000)
        pbh, vbh = (pm, vel) if nbh == n else plummer(nbh)
        if want("bh_1M"):
            configs["bh_1M"] = bh_record(
                ctx, "BASELINE config 4: synthetic %d-body Plummer sphere, Barnes-Hut THETA=0.5, FP64" % nbh,
                pbh, vbh, accuracy=True, cpu_baseline=(world == 1 and not args.no_cpu_baseline))
        if want("bh_1M_uniform"):
            configs["bh_1M_uniform"] = bh_record(
                ctx, "synthetic %d-body uniform cube, Barnes-Hut THETA=0.5, FP64" % nbh,
                pu[:nbh], vu[:nbh], accuracy=False, single_gpu_check=False)
        del pu, vu
        if (world == 8 or args.bh16m) and want("bh_16M"):
            p16, v16 = plummer(16_000_000)
            configs["bh_16M"] = bh_record(
                ctx, "BASELINE config 5: synthetic 16M-body Plummer sphere, Barnes-Hut THETA=0.5, FP64, "
                     "i-slices over %d GPU(s)" % world, p16, v16, accuracy=False)
            del p16, v16
        if world == 1 and want("small_n"):
            configs["small_n"] = small_n_record(ctx)
        for k in ("bh_1M", "bh_1M_uniform", "bh_16M"):
            if k in configs and not bh_parity_ok(configs[k]):
                ok = False

    if rank == 0:
        roof = direct_roofline(nb, prec, uni, my_rows, st0.n_padded, force_ms / args.steps, clocks)
        # one ncu --set full capture of this very launch (1M bodies, FP32, one GPU):
        # dram__bytes_read.sum + dram__bytes_write.sum (profiles/r2_direct_f32_ip_ncu_full.csv,
        # the shipped 256-quad-tile kernel: 28.94 MB + 1.52 MB); other shapes were not captured
        roof["traffic"] = 30467072 if (n == 1_000_000 and world == 1 and prec == nb.PREC_FP32 and uni) else None
        roof["traffic_note"] = ("bytes per launch from ncu (dram read + write); the algorithmic 64 B/body "
                                "include the 32 MB of writes, which leave L2 only after the launch. "
                                "The kernel is FMA-bound: HBM traffic is ~1e-4 of the time budget")
        line = {
            "metric": "pairwise interactions/s", "value": value, "unit": "interactions/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms / args.steps, "steps_per_s": args.steps / (dev_ms * 1e-3),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32" if prec == nb.PREC_FP32 else "f64", "data": "synthetic",
            "config": workload_config(args),
            "sharding": "i-slices over %d GPU(s), positions all-gathered over NVLink every step" % world,
            "clocks": clocks, "wall_s_timed_region": wall,
            "e2e": {"value": e2e_value, "unit": "interactions/s",
                    "h2d_bytes_per_step": nin, "d2h_bytes_per_step": nout,
                    "steps": e2e_steps, "interactions_per_step": 2.0 * n * n,
                    "what": "nb_upload_slice + nb_step(1) + nb_gather_slice(pos, vel) per step, pinned host "
                            "buffers; bytes summed over ranks (each rank moves its own rows over PCIe, "
                            "the ranks exchange the rest over NVLink)",
                    "checksum": checksum},
            "gpu_launches": int(launches),
            "back_to_back": {"ms_per_step": b2b_ms, "value": float(n) * float(n) / (b2b_ms * 1e-3),
                             "split_evaluations": int(stb.split_evaluations),
                             "collectives": int(stb.collectives)},
            "roofline": roof,
            "grid": st0.grid, "block": st0.block,
            "kernel": kernel_name(nb, prec, uni, st0.kernel_variant),
            "parity": parity,
            "configs": configs,
        }
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            rate, kind, sec = cpu_reference_rate(pm, args.cpu_sample, 2, threads)
            line["cpu_baseline"] = {
                "value": rate, "unit": "interactions/s", "cores": threads, "kind": kind,
                "sample": "acceleration() of openmp/exhaustive.c (-O3, FP64) on the first %d bodies "
                          "of the workload, 2 passes, %.1f s" % (args.cpu_sample, sec)}
            line["cpu_baseline"].update(cpu_program_baselines(pm, vel, min(n, 16384), threads))
        if world == 1 and not args.no_ref_gpu:
            f64 = configs.get("fp64_100k") or {}
            bh = configs.get("bh_1M") or {}
            nbh = min(n, 1_000_000)
            p100, v100 = fixture_like(100_000)
            line["ref_gpu_baseline"] = ref_gpu_baseline(
                p100, v100, pm[:nbh] if args.ref_gpu_bh else None, vel[:nbh] if args.ref_gpu_bh else None,
                f64.get("value"), bh.get("ms_per_step"))
        print(json.dumps(line), flush=True)
    ok_all = ctx.reduce([1.0 if ok else 0.0], "min")[0] == 1.0
    ctx.close()
    if not ok_all:
        sys.stderr.write("bench.py: PARITY FAILED (see the \"parity\" records of the JSON line)\n")
        sys.exit(1)


# ------------------------------------------------------- Barnes-Hut arm ----
def run_bh_arm(args):
    """--algorithm bh: one Barnes-Hut config as the whole line (configs 4/5 on their own)."""
    ctx = Ctx(args)
    n = args.bodies
    pm, vel = make_workload(args.distribution, n)
    rec = bh_record(ctx, "synthetic %d-body %s, Barnes-Hut THETA=0.5, fp64" % (n, args.distribution),
                    pm, vel, accuracy=not args.no_accuracy,
                    cpu_baseline=(ctx.world == 1 and not args.no_cpu_baseline))
    ok = bh_parity_ok(rec)
    if ctx.rank == 0:
        line = {"metric": "Barnes-Hut body force evaluations/s", "value": rec["value"], "unit": "bodies/s",
                "n_gpus": ctx.world, "steps": rec["steps"], "warmup": 3, "ms_per_step": rec["ms_per_step"],
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic",
                "config": {"workload": rec["workload"], "bodies": n, "theta": 0.5,
                           "generator": "%s seed 17" % args.distribution,
                           "l2": "flushed between timed steps (256 MiB write)"}}
        for k in ("clocks", "tree", "accuracy_vs_direct", "e2e", "gpu_launches", "roofline", "parity",
                  "cpu_baseline", "cpu_baseline_as_shipped"):
            if k in rec:
                line[k] = rec[k]
        print(json.dumps(line), flush=True)
    ok_all = ctx.reduce([1.0 if ok else 0.0], "min")[0] == 1.0
    ctx.close()
    if not ok_all:
        sys.exit(1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--bodies", type=int, default=1_000_000)
    ap.add_argument("--precision", default="fp32", choices=["fp32", "fp64"])
    ap.add_argument("--distribution", default="plummer", choices=["plummer", "uniform"],
                    help="plummer: Aarseth-Henon-Wielen sphere; uniform: U(-1,1)^3, v = 0 (SURVEY 8d)")
    ap.add_argument("--cpu-sample", type=int, default=32768)
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--ctas-per-sm", type=int, default=0)
    ap.add_argument("--tune-flags", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="headline only: skip the `configs` records")
    ap.add_argument("--only", default="", help="comma list of `configs` records to run (default: all)")
    ap.add_argument("--no-ref-gpu", action="store_true", help="skip the reference's own CUDA programs")
    ap.add_argument("--ref-gpu-bh", action="store_true",
                    help="also try the reference's barnes-hut-cuda at 1M bodies (it does not terminate on "
                         "a B200; run under a 240 s timeout)")
    ap.add_argument("--bh16m", action="store_true", help="measure config 5 (16M-body Barnes-Hut) at any N")
    ap.add_argument("--algorithm", default="direct", choices=["direct", "bh"],
                    help="direct = the headline exhaustive FP32 step; bh = Barnes-Hut (configs 4/5)")
    ap.add_argument("--no-accuracy", action="store_true")
    ap.add_argument("--masses", default="equal", choices=["equal", "spread"],
                    help="equal: m = 1/N (the Plummer model; the FP32 kernel hoists the mass "
                         "multiply); spread: m ~ U(0.5, 1.5)/N, the general kernel")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    elif args.algorithm == "bh":
        run_bh_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
