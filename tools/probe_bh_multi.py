#!/usr/bin/env python
"""Barnes-Hut over W ranks (one host thread per rank): ms per step for the device orders.
    python tools/probe_bh_multi.py WORLD N lopsided|normal [tune_flags,...]
flags: 0 = Morton curve dealt in 32 pieces per rank (default), 64 = contiguous pieces of the
curve, 32 = the caller's order, bits 16..23 = log2(pieces per rank) + 1."""
import os, sys, threading, json
import numpy as np
sys.path.insert(0, "tests")
from conftest import load_package
nb = load_package()
world = int(sys.argv[1]); n = int(sys.argv[2]); kind = sys.argv[3]
rng = np.random.default_rng(5)
if kind == "lopsided":   # a tight, heavy cluster off-centre inside a wide, light halo
    h = n // 2
    a = rng.standard_normal((h, 3)) * 0.05 + np.array([3.0, 2.0, 4.0])
    b = rng.standard_normal((n - h, 3)) * 3.0
    x = np.concatenate([a, b]); rng.shuffle(x)
else:
    x = rng.standard_normal((n, 3))
pm = np.concatenate([x, np.full((n, 1), 1.0 / n)], axis=1)
vel = np.zeros((n, 3))
for flags in ([int(x, 0) for x in sys.argv[4].split(',')] if len(sys.argv) > 4 else (0, nb.TUNE_NO_INTERLEAVE, nb.TUNE_NO_REORDER)):
    cid = nb.comm_id()
    out = [None] * world
    bar = threading.Barrier(world)
    def work(r):
        with nb.Engine(pm, vel, precision=nb.PREC_FP64, algorithm=nb.ALGO_BARNES_HUT, device=r, rank=r,
                       world=world, comm_id_bytes=cid, tune_flags=flags, dt=1e-4) as e:
            e.step(2); e.sync(); bar.wait()
            e.step(4); e.sync()
            out[r] = e.stats().last_step_ms / 4
    ts = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    [t.start() for t in ts]; [t.join() for t in ts]
    print(json.dumps({"kind": kind, "n": n, "world": world, "flags": flags, "ms_per_rank": [round(o, 3) for o in out]}), flush=True)
