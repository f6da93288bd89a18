#!/usr/bin/env python
"""Multi-rank timing probe (GPU box, >= 2 GPUs): ms/step of K back-to-back steps with the
all-gather overlapped (own-slice launch first) and serial (NB_TUNE_NO_OVERLAP).
    python tools/probe_multi.py WORLD PREC N [STEPS]
One host thread per rank; device time from the engines' CUDA events (max over ranks)."""
import json
import os
import sys
import threading

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
from conftest import load_package  # noqa: E402

nb = load_package()


def run(world, prec, n, steps, flags):
    rng = np.random.default_rng(17)
    pm = np.concatenate([rng.standard_normal((n, 3)), np.full((n, 1), 1.0 / n)], axis=1)
    vel = np.zeros((n, 3))
    cid = nb.comm_id() if world > 1 else None
    out = [None] * world
    bar = threading.Barrier(world)

    def work(r):
        with nb.Engine(pm, vel, precision=prec, device=r, rank=r, world=world, comm_id_bytes=cid,
                       tune_flags=flags) as e:
            e.step(2)
            e.sync()
            bar.wait()
            e.step(steps)
            st = e.stats()
            out[r] = (st.last_step_ms / steps, st.split_evaluations, st.collectives)

    ts = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    ms = max(o[0] for o in out)
    print(json.dumps({"world": world, "n": n, "flags": flags, "ms_per_step": ms,
                      "interactions_per_s": float(n) * n / (ms * 1e-3),
                      "split_evaluations": out[0][1], "collectives": out[0][2]}), flush=True)


if __name__ == "__main__":
    world, prec, n = int(sys.argv[1]), sys.argv[2], int(sys.argv[3])
    steps = int(sys.argv[4]) if len(sys.argv) > 4 else 5
    P = nb.PREC_FP32 if prec == "fp32" else nb.PREC_FP64
    for flags in (0, nb.TUNE_NO_OVERLAP, 0, nb.TUNE_NO_OVERLAP):
        run(world, P, n, steps, flags)
