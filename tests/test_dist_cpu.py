"""Host-side logic of the multi-GPU path on CPU: two processes over gloo
(world_size 2, 127.0.0.1).  No engine runs here (there is no CPU engine); what is
exercised is everything the host does around it -- the slice ownership rule
(nbody.slice_bounds == the engine's), the communicator-id hand-off over
torch.distributed, and the step protocol "own slice advances, all-gather, next
evaluation", emulated with the oracle on each rank's slice: the assembled result
must equal the single-process oracle bit for bit.
"""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, steps, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    import oracle as O
    from conftest import golden, load_package
    nb = load_package()
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank,
                            world_size=world)
    try:
        # 1. communicator id: created on rank 0 only, identical everywhere afterwards
        cid = nb.share_comm_id(dist, rank, make=lambda: bytes(range(128)))
        assert cid == bytes(range(128)) and len(cid) == nb.COMM_ID_BYTES
        # 2. slices
        g = golden("1k_bodies")
        pm, vel = g["pos_xyzm"][:n].copy(), g["vel"][:n].copy()
        b, e = nb.slice_bounds(n, world, rank, nb.PREC_FP64)
        bounds = [None] * world
        dist.all_gather_object(bounds, (b, e))
        assert bounds[0][0] == 0 and bounds[-1][1] == n
        assert all(bounds[r][1] == bounds[r + 1][0] for r in range(world - 1))
        # 3. step protocol with the oracle standing in for the kernel on this rank's slice
        lib = O.load_oracle()
        dt = float(np.float32(g["dt"]))

        def acc_slice():
            a = np.zeros((max(e - b, 1), 3))
            if e > b:
                lib.orc_acceleration_slice(n, np.ascontiguousarray(pm), b, e, a)
            return a[:e - b]

        a = acc_slice()
        for _ in range(steps):
            vel[b:e] += a * dt / 2.0
            pm[b:e, :3] += vel[b:e] * dt
            parts = [None] * world
            dist.all_gather_object(parts, pm[b:e].copy())     # the all-gather of the fresh slices
            pm = np.concatenate(parts, axis=0)
            a = acc_slice()
            vel[b:e] += a * dt / 2.0
        vparts = [None] * world
        dist.all_gather_object(vparts, vel[b:e].copy())
        q.put((rank, pm, np.concatenate(vparts, axis=0)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n,steps", [(1000, 3), (257, 2), (100, 1)])
def test_two_process_gloo_step_protocol(n, steps):
    import torch.multiprocessing as mp
    import oracle as O
    from conftest import golden
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, steps, q)) for r in range(2)]
    for p in procs:
        p.start()
    outs = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    g = golden("1k_bodies")
    rp, rv = O.propagate(O.load_oracle(), g["pos_xyzm"][:n], g["vel"][:n], steps, g["dt"])
    for _, pm, v in outs:
        assert np.array_equal(pm, rp) and np.array_equal(v, rv)
