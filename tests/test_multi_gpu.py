"""Multi-GPU path: i-slices over `world` ranks, positions
re-assembled every step by an NCCL all-gather.  One host thread per rank (the
C ABI's threading contract), one engine per device.  These tests need >= 2
B200s on the box (`gpurun --gpus 2`); on a single-GPU box they skip.

Tolerances: a W-rank run sums each body's j-terms in a different order than
the 1-rank run (own slice first, then the gathered slices), so the comparison
is to rounding, not bitwise: FP64 1e-13, FP32 2e-6 relative per body.
"""
import threading

import numpy as np
import pytest

from conftest import golden, rel_err

pytestmark = pytest.mark.gpu


def run_ranks(nb, world, fn, pm, vel, **kw):
    """Runs fn(engine, rank) on `world` engines, one thread and one device each."""
    cid = nb.comm_id()
    out, errs = [None] * world, []

    def work(r):
        try:
            with nb.Engine(pm, vel, device=r, rank=r, world=world, comm_id_bytes=cid, **kw) as e:
                out[r] = fn(e, r)
        except Exception as ex:  # surface in the main thread
            errs.append((r, ex))

    ts = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=300)
    assert not errs, errs
    assert all(o is not None for o in out)
    return out


def need(nb, k):
    if nb.device_count() < k:
        pytest.skip("needs %d GPUs" % k)


@pytest.mark.parametrize("prec", ["fp64", "fp32"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_ranks_agree_with_single_gpu(nb, prec, world):
    need(nb, world)
    g = golden("10k_bodies_g1")
    P = nb.PREC_FP64 if prec == "fp64" else nb.PREC_FP32
    pm, vel = g["pos_xyzm"], g["vel"]
    with nb.Engine(pm, vel, precision=P, dt=g["dt"]) as e:
        a1 = e.accelerations()
        e.step(5)
        p1, v1 = e.gather()

    def fn(e, r):
        acc = e.accelerations()
        e.step(5)
        p, v = e.gather()
        st = e.stats()
        return acc, p, v, (st.slice_begin, st.slice_end, st.force_launches, st.split_evaluations)

    outs = run_ranks(nb, world, fn, pm, vel, precision=P, dt=g["dt"])
    tol = 1e-13 if prec == "fp64" else 2e-6
    covered = 0
    for r, (acc, p, v, (b, e_, launches, split)) in enumerate(outs):
        assert (b, e_) == nb.slice_bounds(len(pm), world, r, P)
        covered += e_ - b
        # accelerations() stores a(x_0), so step 1 is a cheap kick+drift: 1 + 4 + 1 evaluations
        assert launches == 6
        # every evaluation that follows a drift starts on the own slice while the all-gather runs
        assert split == 5
        assert rel_err(acc, a1).max() <= tol
        # 5 steps of the violent 10k fixture amplify rounding by ~1e3 (see test_gpu_parity)
        assert rel_err(p, p1).max() <= tol * 1e3 and rel_err(v, v1).max() <= tol * 1e3
        # every rank holds the same assembled state, bit for bit
        assert np.array_equal(p, outs[0][1]) and np.array_equal(v, outs[0][2])
    assert covered == len(pm)


def test_two_ranks_match_reference_golden(nb):
    """2-rank FP64 run of the 1k fixture against the reference's own numbers."""
    need(nb, 2)
    g = golden("1k_bodies")

    def fn(e, r):
        e.step(10)
        return e.gather()

    outs = run_ranks(nb, 2, fn, g["pos_xyzm"], g["vel"], precision=nb.PREC_FP64, dt=g["dt"])
    for p, v in outs:
        assert rel_err(p, g["pos_10"]).max() <= 1e-12
        assert rel_err(v, g["vel_10"]).max() <= 1e-12


@pytest.mark.parametrize("n", [5, 63, 257, 1000])
def test_ragged_slices(nb, n):
    """slices that end in padding, or are entirely padding (n < world * unit)"""
    need(nb, 2)
    import oracle as O
    rng = np.random.default_rng(n)
    pm = np.concatenate([rng.standard_normal((n, 3)), rng.uniform(0.1, 1, (n, 1))], axis=1)
    vel = rng.standard_normal((n, 3))
    lib = O.load_oracle()
    rp, rv = O.propagate(lib, pm, vel, 3, 0.01)

    def fn(e, r):
        e.step(3)
        return e.gather()

    for p, v in run_ranks(nb, 2, fn, pm, vel, precision=nb.PREC_FP64):
        assert rel_err(p, rp).max() <= 1e-12 and rel_err(v, rv).max() <= 1e-12


def test_energy_and_stepping_patterns_two_ranks(nb):
    need(nb, 2)
    g = golden("1k_bodies")

    def fn(e, r):
        ke0, pe0 = e.energy()
        for _ in range(3):
            e.step(1)
            e.gather()          # closes the step: exercises the stored-acceleration path
        p, v = e.gather()
        return ke0, pe0, p, v

    outs = run_ranks(nb, 2, fn, g["pos_xyzm"], g["vel"], precision=nb.PREC_FP64, dt=g["dt"])
    import oracle as O
    rp, rv = O.propagate(O.load_oracle(), g["pos_xyzm"], g["vel"], 3, g["dt"])
    for ke0, pe0, p, v in outs:
        assert abs(ke0 - g["ke_0"]) <= 1e-12 * abs(g["ke_0"])
        assert abs(pe0 - g["pe_0"]) <= 1e-12 * abs(g["pe_0"])
        assert rel_err(p, rp).max() <= 1e-12 and rel_err(v, rv).max() <= 1e-12


def test_overlap_and_serial_exchange_agree(nb):
    """own-slice-first (two launches around the all-gather) vs wait-then-one-launch"""
    need(nb, 2)
    g = golden("10k_bodies_g1")

    def fn(e, r):
        e.step(3)
        p, v = e.gather()
        return p, v, e.stats().split_evaluations

    a = run_ranks(nb, 2, fn, g["pos_xyzm"], g["vel"], precision=nb.PREC_FP64, dt=g["dt"])
    b = run_ranks(nb, 2, fn, g["pos_xyzm"], g["vel"], precision=nb.PREC_FP64, dt=g["dt"],
                  tune_flags=nb.TUNE_NO_OVERLAP)
    assert a[0][2] == 3 and b[0][2] == 0
    assert rel_err(a[0][0], b[0][0]).max() <= 1e-11 and rel_err(a[0][1], b[0][1]).max() <= 1e-11


@pytest.mark.parametrize("world", [2, 3, 4, 8])
def test_barnes_hut_ranks_bitwise(nb, world):
    """Barnes-Hut over W ranks: every rank builds the same tree from the gathered positions and
    walks the bodies of its own slice; a body's interaction list does not depend on who walks
    it, so the W-rank state equals the 1-rank state bit for bit."""
    need(nb, world)
    g = golden("10k_bodies_g1")
    kw = dict(precision=nb.PREC_FP64, algorithm=nb.ALGO_BARNES_HUT, dt=g["dt"])
    with nb.Engine(g["pos_xyzm"], g["vel"], **kw) as e:
        a1 = e.accelerations()
        e.step(3)
        p1, v1 = e.gather()

    def fn(e, r):
        acc = e.accelerations()
        e.step(3)
        p, v = e.gather()
        return acc, p, v

    for acc, p, v in run_ranks(nb, world, fn, g["pos_xyzm"], g["vel"], **kw):
        assert np.array_equal(acc, a1) and np.array_equal(p, p1) and np.array_equal(v, v1)
    # round 1's replicated sort (every rank sorts all keys) gives the same bits
    for acc, p, v in run_ranks(nb, world, fn, g["pos_xyzm"], g["vel"], tune_flags=nb.TUNE_REPLICATED_SORT, **kw):
        assert np.array_equal(acc, a1) and np.array_equal(p, p1) and np.array_equal(v, v1)


@pytest.mark.parametrize("world,n,every", [(2, 10_000, 2), (2, 1000, 1), (2, 130, 3), (4, 10_000, 3),
                                           (8, 10_000, 2)])
def test_barnes_hut_relayout_keeps_every_bit(nb, monkeypatch, world, n, every):
    """A multi-rank Barnes-Hut engine lays its state out again every K steps (bodies dealt to the
    ranks along the Morton curve of the moment; 32 by default, forced to 1..3 here).  Nothing but
    the place of a body on the devices may change: the state after 7 steps, stepped in one call and
    in pieces, gathered whole and by slices, equals the one-rank state bit for bit."""
    need(nb, world)
    g = golden("10k_bodies_g1")
    pm, vel = g["pos_xyzm"][:n].copy(), g["vel"][:n].copy()
    kw = dict(precision=nb.PREC_FP64, algorithm=nb.ALGO_BARNES_HUT, dt=g["dt"])
    with nb.Engine(pm, vel, **kw) as e:
        e.step(7)
        p1, v1 = e.gather()
        ke1, pe1 = e.energy()
    monkeypatch.setenv("NB_BH_RELAYOUT_EVERY", str(every))

    def fn(e, r):
        e.step(4)
        e.step(1)
        e.step(2)
        b, c = e.io_slice()
        ps, vs = e.gather_slice()
        p, v = e.gather()
        return p, v, ps, vs, b, c, e.energy()

    for p, v, ps, vs, b, c, (ke, pe) in run_ranks(nb, world, fn, pm, vel, **kw):
        assert np.array_equal(p, p1) and np.array_equal(v, v1)
        assert np.array_equal(ps, p1[b:b + c]) and np.array_equal(vs, v1[b:b + c])
        assert abs(ke - ke1) <= 1e-12 * abs(ke1) and abs(pe - pe1) <= 1e-12 * abs(pe1)


@pytest.mark.parametrize("n", [2, 3, 129, 257, 1000])
@pytest.mark.parametrize("flags", ["deal", "curve", "caller"])
def test_barnes_hut_ragged_slices_and_orders(nb, n, flags):
    """Two ranks, sizes around the 128-body chunks the Morton curve is dealt out in (slices that
    are partly or wholly padding), and the three device orders: chunks dealt round robin
    (default), contiguous pieces of the curve, the caller's order.  All equal the 1-rank state."""
    need(nb, 2)
    g = golden("10k_bodies_g1")
    pm, vel = g["pos_xyzm"][:n].copy(), g["vel"][:n].copy()
    tf = {"deal": 0, "curve": nb.TUNE_NO_INTERLEAVE, "caller": nb.TUNE_NO_REORDER}[flags]
    kw = dict(precision=nb.PREC_FP64, algorithm=nb.ALGO_BARNES_HUT, dt=g["dt"])
    with nb.Engine(pm, vel, **kw) as e:
        e.step(2)
        p1, v1 = e.gather()

    def fn(e, r):
        e.step(2)
        return e.gather()

    for p, v in run_ranks(nb, 2, fn, pm, vel, tune_flags=tf, **kw):
        assert np.array_equal(p, p1) and np.array_equal(v, v1)


@pytest.mark.parametrize("algo", ["direct32", "direct64", "bh"])
@pytest.mark.parametrize("world,n", [(2, 1000), (2, 5), (4, 10_000), (8, 10_000)])
def test_slice_io_equals_full_io(nb, algo, world, n):
    """nb_upload_slice / nb_gather_slice: every rank moves only its rows over PCIe, the rest
    travels between the GPUs.  The state is the one nb_upload of all rows gives, bit for bit,
    and the gathered slices concatenate to nb_gather's arrays."""
    need(nb, world)
    g = golden("10k_bodies_g1")
    pm, vel = g["pos_xyzm"][:n].copy(), g["vel"][:n].copy()
    kw = {"direct32": dict(precision=nb.PREC_FP32), "direct64": dict(precision=nb.PREC_FP64),
          "bh": dict(precision=nb.PREC_FP64, algorithm=nb.ALGO_BARNES_HUT)}[algo]
    pm2 = pm.copy()
    pm2[:, :3] *= 1.5                                  # a second state, uploaded by slices
    pm2[:, 3] = np.linspace(0.1, 0.3, n)

    def fn(e, r):
        e.step(2)
        full = e.gather()
        b, c = e.io_slice()
        e.upload_slice(pm2[b:b + c], vel[b:b + c])
        e.step(2)
        part = e.gather_slice()
        e.upload(pm2, vel)
        e.step(2)
        return full, (b, c), part, e.gather()

    outs = run_ranks(nb, world, fn, pm, vel, dt=g["dt"], **kw)
    per = (n + world - 1) // world
    pos_cat = np.concatenate([o[2][0] for o in outs])
    vel_cat = np.concatenate([o[2][1] for o in outs])
    for r, (full, (b, c), part, again) in enumerate(outs):
        assert b == min(n, r * per) and c == min(n - b, per)
        assert np.array_equal(pos_cat, again[0]) and np.array_equal(vel_cat, again[1])
        assert np.array_equal(full[0], outs[0][0][0])
