#!/usr/bin/env python
"""Benchmark of the direct-sum (exhaustive) step -- BASELINE.json's metric:
pairwise interactions/s (and steps/s) at 1M bodies, FP32, on 1/2/4/8 B200.

    python bench.py --gpus N --steps K --warmup W            (N > 1: under torchrun)
    python bench.py --impl reference ...                     (the reference's CPU code)

One "step" = one fused force + leapfrog launch over all N^2 ordered pairs
(incl. j = i, as the reference counts them).  Prints ONE JSON line on rank 0.
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
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)

FLOP_PER_INTERACTION = 20.0   # BASELINE.md section 3 convention
SM_COUNT, FP32_LANES, SM_MAX_MHZ = 148, 128, 1965.0
FP32_PEAK_TFLOPS = SM_COUNT * FP32_LANES * 2 * SM_MAX_MHZ * 1e6 / 1e12   # 74.4
FP64_PEAK_TFLOPS = FP32_PEAK_TFLOPS / 2


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


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.t.join(timeout=2)
        mhz, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                mhz.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, val in zip(names, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(mhz)) if mhz else None,
                "sm_max_mhz": max(mx) if mx else None,
                "power_w": float(np.median(pw)) if pw else None,
                "samples": len(mhz), "reasons": sorted(reasons)}


# ------------------------------------------------------------- CPU legs ----
def cpu_reference_rate(pm, sample_n, evals, threads):
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
    fn()  # warm-up (page faults, thread pool)
    t0 = time.perf_counter()
    for _ in range(evals):
        fn()
    sec = time.perf_counter() - t0
    return evals * float(sample_n) ** 2 / sec, kind, sec


def run_reference_arm(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    pm, _ = plummer(args.bodies)
    sample_n = args.cpu_sample
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
    for _ in range(args.warmup):
        fn()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fn()
    sec = time.perf_counter() - t0
    value = args.steps * float(sample_n) ** 2 / sec
    sample = ("acceleration() of openmp/exhaustive.c (-O3, FP64) on the first %d bodies of the "
              "workload, %d passes" % (sample_n, args.steps))
    line = {
        "impl": "reference", "metric": "pairwise interactions/s", "value": value,
        "unit": "interactions/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": sec / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, 1),
        "cpu_baseline": {"value": value, "unit": "interactions/s", "cores": threads, "kind": kind,
                         "sample": sample},
        "e2e": {"value": value, "unit": "interactions/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, world):
    return {"workload": "synthetic %d-body Plummer sphere, exhaustive direct sum, %s, eps2=0.01, "
                        "dt=0.01f, i-slices over %d GPU(s)" % (args.bodies, args.precision, world),
            "bodies": args.bodies, "precision": args.precision, "generator": "plummer seed 17",
            "l2": "flushed between timed steps (256 MiB write); positions are rewritten every step"}


# -------------------------------------------------------------- GPU arm ----
def run_gpu_arm(args):
    import torch
    from conftest import load_package
    nb = load_package()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        args.gpus = world
    torch.cuda.set_device(local)
    dist = None
    comm = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        # rank 0 creates the engine communicator id; torch.distributed carries it
        comm = nb.share_comm_id(dist, rank, device="cuda")

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    prec = nb.PREC_FP32 if args.precision == "fp32" else nb.PREC_FP64
    n = args.bodies
    pm, vel = plummer(n)
    if args.masses == "spread":
        pm[:, 3] = np.random.default_rng(18).uniform(0.5, 1.5, n) / n
    eng = nb.Engine(pm, vel, precision=prec, device=local, rank=rank, world=world,
                    comm_id_bytes=comm, ctas_per_sm=args.ctas_per_sm, tune_flags=args.tune_flags)
    st0 = eng.stats()
    my_rows = st0.slice_end - st0.slice_begin
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    # ---- kernel-resident throughput: inputs already in HBM -------------------------------
    for _ in range(args.warmup):
        eng.step(1)
    eng.sync()
    sampler = ClockSampler(local)
    launches0 = eng.stats().kernel_launches
    barrier()
    if rank == 0:
        sampler.start()
    dev_ms, force_ms = 0.0, 0.0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        flush.zero_()                 # L2 flush: 256 MiB > 126 MB L2
        torch.cuda.synchronize()
        eng.step(1)
        st = eng.stats()              # syncs the engine stream, reads its CUDA events
        dev_ms += st.last_step_ms
        force_ms += st.last_force_ms
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None
    launches = eng.stats().kernel_launches - launches0
    t = torch.tensor([dev_ms, force_ms], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, force_ms = float(t[0]), float(t[1])
    total_interactions = float(args.steps) * float(n) * float(n)
    value = total_interactions / (dev_ms * 1e-3)

    # ---- the same K steps enqueued back to back (one nb_step call, no L2 flush, no host sync
    # in between): what a simulation run sees; with world > 1 the all-gather of step k overlaps
    # the own-slice launch of step k+1 here.  Reported beside `value`, not instead of it.
    barrier()
    eng.step(args.steps)
    stb = eng.stats()
    tb = torch.tensor([stb.last_step_ms], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(tb, op=dist.ReduceOp.MAX)
    b2b_ms = float(tb[0]) / args.steps

    # ---- end to end through the C ABI with host buffers -----------------------------------
    # one e2e step = propagation() of one time step from host arrays: H2D of positions and
    # velocities (pinned), kick+drift launch, closing force launch, D2H of the new state;
    # it evaluates 2 N^2 interactions, as the reference does for n_steps = 1.
    hp = torch.from_numpy(pm).pin_memory()
    hv = torch.from_numpy(vel).pin_memory()
    op = torch.empty_like(hp).pin_memory()
    ov = torch.empty_like(hv).pin_memory()
    e2e_steps = max(1, min(args.steps, args.e2e_steps))

    def e2e_once():
        eng.upload_raw(hp.data_ptr(), hv.data_ptr())
        eng.step(1)
        eng.gather_raw(op.data_ptr(), ov.data_ptr())

    e2e_once()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_once()
    barrier()
    e2e_wall = time.perf_counter() - t0
    tw = torch.tensor([e2e_wall], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(tw, op=dist.ReduceOp.MAX)
    e2e_value = e2e_steps * 2.0 * float(n) * float(n) / float(tw[0])
    checksum = float(op[:, :3].abs().sum())

    if rank == 0:
        peak = FP32_PEAK_TFLOPS if prec == nb.PREC_FP32 else FP64_PEAK_TFLOPS
        # dominant kernel = the fused force launch; algorithmic flops per launch = 20 * rows * N
        per_launch_flops = FLOP_PER_INTERACTION * float(my_rows) * float(st0.n_padded)
        ach = per_launch_flops / (force_ms / args.steps * 1e-3) / 1e12
        roof = {"bound": "fma", "achieved": ach, "peak": peak, "unit": "TFLOP/s",
                "frac": ach / peak,
                # one ncu --set full capture of this very launch (1M bodies, FP32, one GPU):
                # dram__bytes_read.sum 29.02 MB + dram__bytes_write.sum 1.65 MB
                # (profiles/r1_f32_direct_1M_ncu_full.csv); other shapes were not captured
                "traffic": 30671104 if (n == 1_000_000 and world == 1 and prec == nb.PREC_FP32
                                        and st0.uniform_mass) else None,
                "peak_source": "nominal: 148 SM x 128 FP32 lanes x 2 x 1965 MHz "
                               "(MEASURED_PEAKS.json has no FP32 FMA figure; sm_max_mhz from it)",
                "flop_per_interaction": FLOP_PER_INTERACTION,
                # tools/microbench.cu on this pool's B200 (profiles/r1_microbench_pipe_rates.jsonl):
                # FFMA 123.1 lane-ops/clk/SM; 6 FFMA2 + 1 MUFU (this kernel's mix) 12.5 clk per
                # interaction-warp -> the mix's ceiling at sm_max_mhz
                "peak_measured_ffma": SM_COUNT * 123.09 * 2 * SM_MAX_MHZ * 1e6 / 1e12,
                "ceiling_of_instruction_mix": (SM_COUNT * 4 * 32 / (12.5 * (11.0 / 12.0 if st0.uniform_mass else 1.0))
                                               * SM_MAX_MHZ * 1e6
                                               * FLOP_PER_INTERACTION / 1e12) if prec == nb.PREC_FP32
                else (SM_COUNT * 62.2 / 16.0 * SM_MAX_MHZ * 1e6 * FLOP_PER_INTERACTION / 1e12),
                "traffic_note": "bytes per launch from ncu (dram read + write); the algorithmic 64 B/body "
                                "include the 32 MB of writes, which leave L2 only after the launch. "
                                "The kernel is FMA-bound: HBM traffic is ~1e-4 of the time budget",
                "hbm_algorithmic_bytes_per_launch": 64.0 * my_rows if prec == nb.PREC_FP32
                else 128.0 * my_rows}
        if clocks and clocks.get("sm_mhz"):
            roof["frac_at_observed_clock"] = ach / (peak * clocks["sm_mhz"] / SM_MAX_MHZ)
        # the HBM side of the same launch, against the pool's measured copy bandwidth: shows by how
        # far the kernel is NOT memory-bound (algorithmic bytes / launch time vs MEASURED_PEAKS.json)
        hbm_peak, hbm_src = 6528.0, "fallback of the profiling recipe"
        try:
            with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "MEASURED_PEAKS.json")) as f:
                hbm_peak, hbm_src = float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
        except Exception:
            pass
        hbm_ach = roof["hbm_algorithmic_bytes_per_launch"] / (force_ms / args.steps * 1e-3) / 1e9
        roof["hbm"] = {"achieved": hbm_ach, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_ach / hbm_peak,
                       "peak_source": hbm_src}
        line = {
            "metric": "pairwise interactions/s", "value": value, "unit": "interactions/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms / args.steps, "steps_per_s": args.steps / (dev_ms * 1e-3),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32" if prec == nb.PREC_FP32 else "f64", "data": "synthetic",
            "config": workload_config(args, world),
            "clocks": clocks, "wall_s_timed_region": wall,
            "e2e": {"value": e2e_value, "unit": "interactions/s",
                    "h2d_bytes_per_step": int(hp.numel() * 8 + hv.numel() * 8),
                    "d2h_bytes_per_step": int(op.numel() * 8 + ov.numel() * 8),
                    "steps": e2e_steps, "interactions_per_step": 2.0 * n * n,
                    "what": "nb_upload + nb_step(1) + nb_gather(pos, vel) per step, pinned host buffers",
                    "checksum": checksum},
            "gpu_launches": int(launches),
            "back_to_back": {"ms_per_step": b2b_ms, "value": float(n) * float(n) / (b2b_ms * 1e-3),
                             "split_evaluations": int(stb.split_evaluations),
                             "collectives": int(stb.collectives)},
            "roofline": roof,
            "grid": st0.grid, "block": st0.block,
            "kernel": ("packed f32x2, equal masses: m_j hoisted out of the loop, 11 FMA-pipe ops + 1 MUFU "
                       "per interaction" if eng.stats().uniform_mass else
                       "packed f32x2, general masses: 12 FMA-pipe ops + 1 MUFU per interaction")
            if prec == nb.PREC_FP32 else "fp64: 16 FP64-pipe ops + 1 MUFU per interaction",
        }
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            rate, kind, sec = cpu_reference_rate(pm, args.cpu_sample, 2, threads)
            line["cpu_baseline"] = {
                "value": rate, "unit": "interactions/s", "cores": threads, "kind": kind,
                "sample": "acceleration() of openmp/exhaustive.c (-O3, FP64) on the first %d bodies "
                          "of the workload, 2 passes, %.1f s" % (args.cpu_sample, sec)}
        print(json.dumps(line), flush=True)
    eng.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


# ------------------------------------------------------- Barnes-Hut arm ----
def bh_cpu_baseline(pm, vel, sample_n, threads):
    """The reference's barnes-hut-omp (openmp/barnes-hut.c, -O3, compiled in place into
    oracle/_ref) on the first sample_n bodies, one step = two force evaluations; timed by its
    own 'Elapsed wall time' line (the CSV read is outside it, as in the reference)."""
    import re
    import tempfile
    import oracle as O
    from conftest import load_package
    exe = os.path.join(O.REFDIR, "barnes_hut_omp_fast")
    if not os.path.exists(exe):
        return None
    h = load_package().load_host_library()
    with tempfile.TemporaryDirectory() as td:
        inp = os.path.join(td, "in.csv")
        h.nbh_write_bodies(inp.encode(), sample_n, np.ascontiguousarray(pm[:sample_n]),
                           np.ascontiguousarray(vel[:sample_n]))
        r = subprocess.run([exe, inp, "0", "0.01", "0.01", os.path.join(td, "o.csv"), str(threads)],
                           capture_output=True, text=True, timeout=600)
    m = re.search(r"Elapsed wall time: ([0-9.]+)", r.stdout)
    if not m:
        return None
    sec = float(m.group(1))
    return {"value": 2.0 * sample_n / sec, "unit": "body force evaluations/s", "cores": threads,
            "kind": "reference",
            "sample": "barnes-hut-omp (openmp/barnes-hut.c, -O3) on the first %d bodies, 1 step "
                      "(2 evaluations), %.2f s by its own clock" % (sample_n, sec)}


def run_bh_arm(args):
    """BASELINE configs 4/5: Barnes-Hut (THETA = 0.5, FP64) on a Plummer sphere; value = body
    force evaluations per second (bodies x steps / time), accuracy against the direct-sum
    kernel on the same positions when that is affordable (<= 2M bodies, 1 GPU)."""
    import torch
    from conftest import load_package, rel_err
    nb = load_package()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist, comm = None, None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        comm = nb.share_comm_id(dist, rank, device="cuda")

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    n = args.bodies
    pm, vel = plummer(n)
    eng = nb.Engine(pm, vel, precision=nb.PREC_FP64, algorithm=nb.ALGO_BARNES_HUT, device=local,
                    rank=rank, world=world, comm_id_bytes=comm, tune_flags=args.tune_flags)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(args.warmup):
        eng.step(1)
    eng.sync()
    l0 = eng.stats().kernel_launches
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    dev_ms = 0.0
    for _ in range(args.steps):
        flush.zero_()
        torch.cuda.synchronize()
        eng.step(1)
        dev_ms += eng.stats().last_step_ms
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = eng.stats().kernel_launches - l0
    t = torch.tensor([dev_ms], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms = float(t[0])
    # end to end through the C ABI: upload, one step, gather (2 evaluations)
    hp, hv = torch.from_numpy(pm).pin_memory(), torch.from_numpy(vel).pin_memory()
    op, ov = torch.empty_like(hp).pin_memory(), torch.empty_like(hv).pin_memory()

    def e2e_once():
        eng.upload_raw(hp.data_ptr(), hv.data_ptr())
        eng.step(1)
        eng.gather_raw(op.data_ptr(), ov.data_ptr())

    e2e_once()
    barrier()
    t0 = time.perf_counter()
    for _ in range(3):
        e2e_once()
    barrier()
    tw = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(tw, op=dist.ReduceOp.MAX)
    e2e_value = 3 * 2.0 * n / float(tw[0])
    acc_stats = None
    if world == 1 and n <= 2_000_000 and not args.no_accuracy:
        eng.upload(pm, vel)
        a_bh = eng.accelerations()
        with nb.Engine(pm, vel, precision=nb.PREC_FP64, device=local) as d:
            a_dir = d.accelerations()
        err = rel_err(a_bh, a_dir)
        acc_stats = {"median": float(np.median(err)), "p99": float(np.percentile(err, 99)),
                     "max": float(err.max()),
                     "what": "per-body relative acceleration error vs the FP64 direct-sum kernel"}
    # tree statistics from a second, untimed engine: the counting walk is ~10 % slower
    eng.close()
    if dist is not None:
        comm = nb.share_comm_id(dist, rank, device="cuda")
    with nb.Engine(pm, vel, precision=nb.PREC_FP64, algorithm=nb.ALGO_BARNES_HUT, device=local,
                   rank=rank, world=world, comm_id_bytes=comm,
                   tune_flags=args.tune_flags | nb.TUNE_COUNT_VISITS) as ce:
        ce.step(1)
        ti = ce.tree_stats()
    if rank == 0:
        my = n / world
        line = {
            "metric": "Barnes-Hut body force evaluations/s", "value": n * args.steps / (dev_ms * 1e-3),
            "unit": "bodies/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "synthetic %d-body Plummer sphere, Barnes-Hut THETA=0.5, fp64, "
                                   "i-slices over %d GPU(s), tree replicated" % (n, world),
                       "bodies": n, "theta": 0.5, "generator": "plummer seed 17",
                       "l2": "flushed between timed steps (256 MiB write)"},
            "clocks": clocks,
            "tree": {"nodes": ti.nodes, "depth": ti.depth,
                     "interactions_per_body": ti.interactions / my, "visits_per_body": ti.node_visits / my},
            "accuracy_vs_direct": acc_stats,
            "e2e": {"value": e2e_value, "unit": "bodies/s", "h2d_bytes_per_step": int(hp.numel() * 8 + hv.numel() * 8),
                    "d2h_bytes_per_step": int(op.numel() * 8 + ov.numel() * 8),
                    "what": "nb_upload + nb_step(1) + nb_gather per step = 2 evaluations"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": None, "peak": None, "unit": "GB/s", "frac": None,
                         "traffic": None, "note": "latency/issue-bound tree walk, see DESIGN.md 4.3"},
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = bh_cpu_baseline(pm, vel, min(n, 200_000), os.cpu_count() or 1)
        print(json.dumps(line), flush=True)
    eng.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--bodies", type=int, default=1_000_000)
    ap.add_argument("--precision", default="fp32", choices=["fp32", "fp64"])
    ap.add_argument("--cpu-sample", type=int, default=32768)
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--ctas-per-sm", type=int, default=0)
    ap.add_argument("--tune-flags", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--algorithm", default="direct", choices=["direct", "bh"],
                    help="direct = the headline exhaustive FP32 step; bh = Barnes-Hut (configs 4/5)")
    ap.add_argument("--no-accuracy", action="store_true")
    ap.add_argument("--masses", default="equal", choices=["equal", "spread"],
                    help="equal: m = 1/N (the Plummer model; the FP32 kernel hoists the mass "
                         "multiply); spread: m ~ U(0.5, 1.5)/N, the general kernel")
    args = ap.parse_args()
    if args.algorithm == "bh" and args.impl != "reference":
        run_bh_arm(args)
    elif args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
