#!/usr/bin/env python
"""Kernel timing probe (GPU box): interactions/s of the direct-sum step for a list of
configurations.  Device time comes from the engine's own CUDA events (nb_get_stats).
    python tools/probe_perf.py fp32:1000000 fp32:1000000:scalar fp64:100000 ...
Spec: prec:N[:opt[,opt]]  opts: scalar, nosplit, ctas=K, variant=V, steps=S
"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
from conftest import load_package  # noqa: E402

nb = load_package()


def run(spec):
    parts = spec.split(":")
    prec = nb.PREC_FP32 if parts[0] == "fp32" else nb.PREC_FP64
    algo = nb.ALGO_BARNES_HUT if parts[0] == "bh" else nb.ALGO_DIRECT
    plummer = False
    n = int(parts[1])
    opts = parts[2].split(",") if len(parts) > 2 else []
    flags, ctas, steps, burst = 0, 0, 3, 0
    for o in opts:
        if o == "scalar":
            flags |= nb.TUNE_SCALAR_FP32
        elif o == "nosplit":
            flags |= nb.TUNE_NO_SPLIT
        elif o.startswith("ctas="):
            ctas = int(o[5:])
        elif o.startswith("variant="):
            flags |= int(o[8:]) << 8
        elif o.startswith("steps="):
            steps = int(o[6:])
        elif o.startswith("emulate="):  # Barnes-Hut on one GPU: sort as K slices + merge (the multi-rank pipeline)
            flags |= int(o[8:]) << nb.TUNE_EMULATE_RUNS_SHIFT
        elif o.startswith("pieces="):  # Barnes-Hut, world > 1: 2^K curve pieces per rank
            flags |= (int(o[7:]) + 1) << 16
        elif o.startswith("burst="):
            burst = int(o[6:])
        elif o == "nouniform":
            flags |= nb.TUNE_NO_UNIFORM
        elif o == "nograph":
            flags |= nb.TUNE_NO_GRAPH
        elif o == "noreorder":
            flags |= nb.TUNE_NO_REORDER
        elif o == "count":
            flags |= nb.TUNE_COUNT_VISITS
        elif o == "plummer":
            plummer = True
    rng = np.random.default_rng(17)
    pm = np.concatenate([rng.standard_normal((n, 3)), np.full((n, 1), 1.0 / n)], axis=1)
    vel = np.zeros((n, 3))
    if plummer:
        sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
        from bench import plummer as gen
        pm, vel = gen(n)
    with nb.Engine(pm, vel, precision=prec, algorithm=algo, ctas_per_sm=ctas, tune_flags=flags) as e:
        e.step(2)  # warm-up
        e.sync()
        times = []
        for _ in range(steps):
            e.step(1)
            st = e.stats()
            times.append(st.last_force_ms)
        st = e.stats()
        if burst: # one nb_step(burst) call: what a caller that does not read back every step sees
            per = []
            for _ in range(5):
                e.step(burst)
                per.append(e.stats().last_step_ms / burst * 1e3)
            print(json.dumps({"spec": spec, "n": n, "us_per_step_in_burst": [round(x, 2) for x in per],
                              "grid": st.grid, "block": st.block, "variant": st.kernel_variant}), flush=True)
            return
        best, med = min(times), float(np.median(times))
        out = {"spec": spec, "n": n, "grid": st.grid, "block": st.block,
               "force_ms_min": best, "force_ms_med": med,
               "interactions_per_s": n * float(st.n_padded) / (med * 1e-3),
               "frac_fp32_roofline_20flop_1965MHz": n * float(st.n_padded) / (med * 1e-3) / 3.72e12}
        if algo == nb.ALGO_BARNES_HUT:
            ti = e.tree_stats()
            out.update({"nodes": ti.nodes, "depth": ti.depth, "interactions_per_body": ti.interactions / n,
                        "visits_per_body": ti.node_visits / n, "launches_per_eval": None,
                        "stage_ms": [round(x, 4) for x in ti.stage_ms[:7]],
                        "lib": os.path.basename(nb.nbody.LIB_PATH) if hasattr(nb, "nbody") else None})
            del out["frac_fp32_roofline_20flop_1965MHz"], out["interactions_per_s"]
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    for s in sys.argv[1:]:
        try:
            run(s)
        except Exception as ex:  # keep going: this is a probe
            print(json.dumps({"spec": s, "error": str(ex)}), flush=True)
