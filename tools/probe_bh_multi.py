#!/usr/bin/env python3
"""Barnes-Hut step anatomy on W ranks (run under torchrun, or plainly for W = 1).

  torchrun --nnodes=1 --nproc-per-node W --master-addr 127.0.0.1 tools/probe_bh_multi.py \
      N[:opt[,opt]] ...
opts: pieces=LOG2 (curve pieces per rank, 2^LOG2), replicated (every rank sorts everything),
      nointerleave, steps=K, uniform

Per spec one JSON line from rank 0: back-to-back ms/step with CUDA graphs, and for an engine with
plain launches the stage times (max and min over ranks), launch_force and nb_step as wholes.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402

NAMES = ["keys", "sort", "merge", "order", "build", "com", "traversal", "exchange"]


def run(ctx, spec):
    nb = ctx.nb
    parts = spec.split(":")
    n = int(parts[0])
    flags, steps, uniform = 0, 20, False
    for o in (parts[1].split(",") if len(parts) > 1 else []):
        if o.startswith("pieces="):
            flags |= (int(o[7:]) + 1) << 16
        elif o == "replicated":
            flags |= nb.TUNE_REPLICATED_SORT
        elif o == "nointerleave":
            flags |= nb.TUNE_NO_INTERLEAVE
        elif o.startswith("steps="):
            steps = int(o[6:])
        elif o == "uniform":
            uniform = True
    pm, vel = bench.uniform_cube(n) if uniform else bench.plummer(n)
    kw = dict(precision=nb.PREC_FP64, algorithm=nb.ALGO_BARNES_HUT)
    out = {"spec": spec, "world": ctx.world, "n": n}
    with ctx.engine(pm, vel, tune_flags=flags, **kw) as eng:
        eng.step(4)
        eng.sync()
        ctx.barrier()
        eng.step(steps)
        out["b2b_ms_per_step"] = ctx.reduce([eng.stats().last_step_ms], "max")[0] / steps
        one = []
        for _ in range(5):
            ctx.barrier()
            eng.step(1)
            one.append(ctx.reduce([eng.stats().last_step_ms], "max")[0])
        out["one_step_ms"] = [round(x, 3) for x in one]
    with ctx.engine(pm, vel, tune_flags=flags | nb.TUNE_NO_GRAPH, **kw) as se:
        se.step(3)
        se.sync()
        ctx.barrier()
        se.step(1)
        st = se.stats()
        whole = ctx.reduce([st.last_force_ms, st.last_step_ms], "max")
        sm = list(se.tree_stats().stage_ms)
        mx, mn = ctx.reduce(sm, "max"), ctx.reduce(sm, "min")
        out["plain"] = {"force_ms": round(whole[0], 3), "step_ms": round(whole[1], 3),
                        "stage_ms_max": {k: round(v, 3) for k, v in zip(NAMES, mx)},
                        "stage_ms_min": {k: round(v, 3) for k, v in zip(NAMES, mn)},
                        "stage_sum_max": round(sum(mx[:7]), 3)}
        ctx.barrier()
        se.step(steps)
        out["plain"]["b2b_ms_per_step"] = ctx.reduce([se.stats().last_step_ms], "max")[0] / steps
    if ctx.rank == 0:
        print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("specs", nargs="+")
    a = ap.parse_args()
    ctx = bench.Ctx(argparse.Namespace())
    try:
        for s in a.specs:
            run(ctx, s)
    finally:
        ctx.close()


if __name__ == "__main__":
    main()
