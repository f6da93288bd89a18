"""GPU parity tests of the Barnes-Hut path:
Morton/radix-sort octree, bottom-up centre of mass, warp-cooperative THETA
traversal, against the reference's serial Barnes-Hut (tests/golden/*.npz holds
bh_acc0 / bh_stats produced by the reference's own compiled sources, and
oracle/bh_oracle.c restates it bit for bit).

Tolerances: the GPU tree has the reference's cells and each body walks the
reference's interaction list in the reference's order, so accelerations agree
to rounding: <= 1e-12 relative per body (the only differences are rsqrt^3 for
pow(x,-1.5) and FMA contraction).  Against the DIRECT sum the bound is the
reference's own THETA = 0.5 envelope: median <= 4e-3,
p99 <= 2e-2, max <= 1.2e-1.
"""
import numpy as np
import pytest

from conftest import golden, rel_err

pytestmark = pytest.mark.gpu

CASES = ["10_bodies", "20_bodies_g1", "33_bodies", "100_bodies_g1", "500_bodies_g1", "1k_bodies",
         "10k_bodies_g1", "sun_earth"]


def bh_engine(nb, pm, vel, **kw):
    return nb.Engine(pm, vel, precision=nb.PREC_FP64, algorithm=nb.ALGO_BARNES_HUT, **kw)


@pytest.mark.parametrize("name", CASES)
def test_bh_accelerations_match_reference(nb, oracle_lib, name):
    g = golden(name)
    with bh_engine(nb, g["pos_xyzm"], g["vel"], tune_flags=nb.TUNE_COUNT_VISITS) as e:
        acc = e.accelerations()
        ti = e.tree_stats()
    n = len(g["pos_xyzm"])
    ost, oacc = np.zeros(5), np.zeros((n, 3))
    oracle_lib.orc_bh_acceleration(n, np.ascontiguousarray(g["pos_xyzm"]), oacc, ost.ctypes.data)
    inter, visits = ost[3], ost[4]                # per body, from the oracle's walk
    if "bh_stats" in g.files:                     # from the reference's own tree
        cells, depth, border = g["bh_stats"]
        ref = g["bh_acc0"]
    else:                                         # sun_earth: no reference Barnes-Hut vector kept
        cells, depth, border = ost[:3]
        ref = oacc
    assert rel_err(acc, ref).max() <= 1e-12
    # the same tree: cell count, depth, root cube; the same walk: interaction and visit totals
    assert ti.nodes == int(cells) and ti.depth == int(depth)
    assert ti.root_border == border
    assert ti.interactions == int(round(inter * n)) and ti.node_visits == int(round(visits * n))


@pytest.mark.parametrize("name,k", [("100_bodies_g1", 10), ("1k_bodies", 10), ("1k_bodies", 100)])
def test_bh_state_after_k_steps(nb, oracle_lib, name, k):
    """KDK loop with the tree rebuilt after every drift (serial/barnes-hut.c:445-517)."""
    g = golden(name)
    pm, v = g["pos_xyzm"].copy(), g["vel"].copy()
    rc = oracle_lib.orc_bh_propagate(len(pm), pm, v, k, np.float32(g["dt"]))
    assert rc == 0
    with bh_engine(nb, g["pos_xyzm"], g["vel"], dt=g["dt"]) as e:
        e.step(k)
        pos, vel = e.gather()
        assert e.stats().force_launches == k + 1
    # an accept/open decision on the knife edge flips with rounding and changes a force by
    # ~1e-3: the discrete tree makes long runs less tight than the direct sum's 1e-10
    tol = 1e-11 if k <= 10 else 1e-8
    assert rel_err(pos, pm).max() <= tol and rel_err(vel, v).max() <= tol


@pytest.mark.parametrize("name", ["1k_bodies", "10k_bodies_g1"])
def test_bh_error_envelope_vs_direct(nb, name):
    g = golden(name)
    with bh_engine(nb, g["pos_xyzm"], g["vel"]) as e:
        acc = e.accelerations()
    err = rel_err(acc, g["acc0"])  # acc0: the reference's direct sum
    assert np.median(err) <= 4e-3 and np.percentile(err, 99) <= 2e-2 and err.max() <= 1.2e-1
    assert np.median(err) > 1e-5  # it IS an approximation: a silent direct sum would pass the bound


@pytest.mark.parametrize("n", [1, 2, 3, 9, 31, 33, 64, 129, 1000, 4097])
def test_bh_ragged_sizes(nb, oracle_lib, n):
    rng = np.random.default_rng(500 + n)
    pm = np.concatenate([rng.standard_normal((n, 3)), rng.uniform(0.1, 1.0, (n, 1))], axis=1)
    vel = rng.standard_normal((n, 3))
    ref = np.zeros((n, 3))
    oracle_lib.orc_bh_acceleration(n, pm, ref, None)
    rp, rv = pm.copy(), vel.copy()
    oracle_lib.orc_bh_propagate(n, rp, rv, 3, np.float32(0.01))
    with bh_engine(nb, pm, vel) as e:
        acc = e.accelerations()
        e.step(3)
        pos, v = e.gather()
    scale = max(np.abs(ref).max(), 1e-30)
    assert np.abs(acc - ref).max() <= 1e-12 * scale
    assert rel_err(pos, rp).max() <= 1e-11 and rel_err(v, rv).max() <= 1e-11


def test_bh_empty(nb):
    with bh_engine(nb, np.zeros((0, 4)), np.zeros((0, 3))) as e:
        e.step(2)
        pos, vel = e.gather()
    assert pos.shape == (0, 4)


def test_bh_theta_zero_is_the_direct_sum(nb):
    """theta = 0 accepts no cell: every leaf is visited, the sum is the exhaustive one."""
    g = golden("500_bodies_g1")
    with bh_engine(nb, g["pos_xyzm"], g["vel"], theta=0.0) as e:
        acc = e.accelerations()
    assert rel_err(acc, g["acc0"]).max() <= 1e-12


def test_bh_coincident_bodies_bucket(nb):
    """Bodies that share a finest (level 21) cell: the reference would subdivide for ever
    (serial/barnes-hut.c:208-238 loops on coincident bodies); here they form a bucket leaf
    that is summed directly.  Checked against the direct-sum engine."""
    rng = np.random.default_rng(3)
    n = 300
    pm = np.concatenate([rng.standard_normal((n, 3)), rng.uniform(0.1, 1.0, (n, 1))], axis=1)
    pm[17, :3] = pm[5, :3]                     # exactly coincident
    pm[40, :3] = pm[41, :3] + 1e-9             # same finest cell, not coincident
    pm[100:104, :3] = pm[99, :3]               # five on one point
    vel = np.zeros((n, 3))
    with nb.Engine(pm, vel, precision=nb.PREC_FP64) as e:
        ref = e.accelerations()
    with bh_engine(nb, pm, vel, theta=0.0) as e:
        acc = e.accelerations()
        ti = e.tree_stats()
    assert ti.depth == 22                      # root .. level 21
    assert rel_err(acc, ref).max() <= 1e-11
    with bh_engine(nb, pm, vel) as e:
        err = rel_err(e.accelerations(), ref)
    assert err.max() <= 1.2e-1 and np.median(err) <= 1e-2


def test_bh_node_pool_exhaustion_is_an_error(nb):
    """Tight pairs spread over the box need a chain of single-child cells each, down to the
    finest level: far more cells than the pool (2n + 4096) holds.  The reference prints
    "Too many nodes" and exits (openmp/barnes-hut.c:158-161); the engine finishes the walk on
    a truncated tree without touching invalid memory and refuses to hand out the result."""
    rng = np.random.default_rng(23)
    half = 4000
    a = rng.uniform(-1, 1, (half, 3))
    pm = np.concatenate([np.concatenate([a, a + 1e-8]), np.ones((2 * half, 1))], axis=1)
    vel = np.zeros((2 * half, 3))
    with bh_engine(nb, pm, vel) as e:
        with pytest.raises(nb.NbodyError) as ei:
            e.accelerations()
            e.gather()
        assert ei.value.code == -3             # NB_E_NOMEM
        assert "pool" in str(ei.value)


def test_bh_overflow_flags_are_cleared_by_a_new_state(nb):
    """The overflow flags stay set between evaluations (a truncated tree is caught at the next
    gather) but a new, harmless state uploaded into the same engine must work again."""
    rng = np.random.default_rng(23)
    half = 4000
    a = rng.uniform(-1, 1, (half, 3))
    pm = np.concatenate([np.concatenate([a, a + 1e-8]), np.ones((2 * half, 1))], axis=1)
    vel = np.zeros((2 * half, 3))
    ok = np.concatenate([rng.standard_normal((2 * half, 3)), np.ones((2 * half, 1))], axis=1)
    with bh_engine(nb, pm, vel) as e:
        with pytest.raises(nb.NbodyError):
            e.accelerations()
            e.gather()
        with pytest.raises(nb.NbodyError):    # still refused: sticky between evaluations
            e.gather()
        e.upload(ok, vel)
        acc = e.accelerations()
        e.gather()
    with nb.Engine(ok, vel, precision=nb.PREC_FP64) as d:
        err = rel_err(acc, d.accelerations())
    assert err.max() <= 1.2e-1 and np.median(err) <= 1e-2


def _off_centre_system():
    """A cluster around (10,10,10) plus one body at (-9,-9,-9): the root cube is centred on the
    ORIGIN (serial/barnes-hut.c:326-338), the centre of mass is far off-centre, and the lone
    body is more than one root border away from it."""
    rng = np.random.default_rng(77)
    n = 400
    pos = 10.0 + 0.3 * rng.standard_normal((n, 3))
    pos[0] = (-9.0, -9.0, -9.0)
    return np.concatenate([pos, rng.uniform(0.1, 1.0, (n, 1))], axis=1)


def _tiny_system():
    """|coord| < 0.05: border^2 of the root is below eps2 = 0.01."""
    rng = np.random.default_rng(78)
    n = 300
    pos = rng.uniform(-0.04, 0.04, (n, 3))
    return np.concatenate([pos, rng.uniform(0.1, 1.0, (n, 1))], axis=1)


@pytest.mark.parametrize("make", [_off_centre_system, _tiny_system])
def test_bh_root_is_tested_like_any_cell(nb, oracle_lib, make):
    """The root goes through the same acceptance test as every cell (border / dist < THETA,
    serial/barnes-hut.c:405,434): with THETA = 0.5 it can never be accepted (both the body and
    the centre of mass lie in the cube), with THETA = 0 the walk is the exact direct sum.
    Catches a root threshold that forgets THETA (round-1 advisor finding)."""
    pm = make()
    n = len(pm)
    vel = np.zeros((n, 3))
    ref = np.zeros((n, 3))
    ost = np.zeros(5)
    oracle_lib.orc_bh_acceleration(n, pm, ref, ost.ctypes.data)
    with bh_engine(nb, pm, vel, tune_flags=nb.TUNE_COUNT_VISITS) as e:
        acc = e.accelerations()
        ti = e.tree_stats()
    scale = np.abs(ref).max()
    assert np.abs(acc - ref).max() <= 1e-12 * scale
    assert ti.interactions == int(round(ost[3] * n)) and ti.node_visits == int(round(ost[4] * n))
    direct = np.zeros((n, 3))
    oracle_lib.orc_acceleration(n, pm, direct)
    with bh_engine(nb, pm, vel, theta=0.0) as e:
        exact = e.accelerations()
    assert np.abs(exact - direct).max() <= 1e-12 * np.abs(direct).max()


def test_bh_fp32_is_refused(nb):
    g = golden("10_bodies")
    with pytest.raises(nb.NbodyError) as ei:
        nb.Engine(g["pos_xyzm"], g["vel"], precision=nb.PREC_FP32, algorithm=nb.ALGO_BARNES_HUT)
    assert ei.value.code == -1


@pytest.mark.parametrize("n", [200_000, 1_000_000])   # 1M: BASELINE config 4
def test_bh_large_n_envelope(nb, n):
    """Barnes-Hut against the direct-sum kernel on the same positions (the "direct-sum accuracy
    check" of BASELINE config 4), plus two properties that need no reference: THETA = 0 opens
    every cell, so a sample of bodies must reproduce the direct sum to rounding; and the counted
    interactions of two engines on the same input agree exactly (deterministic lists)."""
    rng = np.random.default_rng(17)
    pm = np.concatenate([rng.standard_normal((n, 3)), np.full((n, 1), 1.0 / n)], axis=1)
    vel = np.zeros((n, 3))
    with nb.Engine(pm, vel, precision=nb.PREC_FP64) as e:
        ref = e.accelerations()
    with bh_engine(nb, pm, vel, tune_flags=nb.TUNE_COUNT_VISITS) as e:
        acc = e.accelerations()
        ti = e.tree_stats()
    err = rel_err(acc, ref)
    assert np.median(err) <= 4e-3 and np.percentile(err, 99) <= 2e-2 and err.max() <= 1.2e-1
    assert 0.3 * n < ti.nodes < 1.0 * n          # the reference measures ~0.49 N internal cells
    assert 500 < ti.interactions / n < 5000
    with bh_engine(nb, pm, vel, tune_flags=nb.TUNE_COUNT_VISITS) as e:
        acc_again = e.accelerations()
        ti2 = e.tree_stats()
    assert np.array_equal(acc_again, acc) and ti2.interactions == ti.interactions
    if n <= 200_000:                             # THETA = 0 is the O(N^2) walk: keep it small
        with bh_engine(nb, pm, vel, theta=0.0) as e:
            exact = e.accelerations()
        assert rel_err(exact, ref).max() <= 1e-10


def test_bh_device_order_is_invisible(nb):
    """The engine keeps the bodies in Morton order on the device (rank slices = compact regions);
    every output comes back in the caller's order, bit-identical to an engine that keeps the
    caller's order (NB_TUNE_NO_REORDER)."""
    g = golden("1k_bodies")
    res = []
    for flags in (0, nb.TUNE_NO_REORDER):
        with bh_engine(nb, g["pos_xyzm"], g["vel"], tune_flags=flags) as e:
            acc = e.accelerations()
            buf = e.frame_buffer()
            e.step(2)
            e.frame_begin(buf)
            e.step(1)
            e.frame_wait(buf)
            frame = buf.copy()
            pos, vel = e.gather()
            ke, pe = e.energy()
        res.append((acc, frame, pos, vel, ke, pe))
    for a, b in zip(res[0][:4], res[1][:4]):
        assert np.array_equal(a, b)
    assert np.array_equal(res[0][2][:, 3], g["pos_xyzm"][:, 3])      # masses where the caller put them
    assert abs(res[0][4] - res[1][4]) <= 1e-12 * abs(res[1][4])      # sums in another order
    assert abs(res[0][5] - res[1][5]) <= 1e-12 * abs(res[1][5])


def test_bh_graph_replay_equals_plain_launches(nb):
    """An evaluation is captured once per (ping-pong side, epilogue flags) as a CUDA graph and
    replayed; stepping patterns that mix all the epilogues give bit-identical states with the
    graphs switched off (NB_TUNE_NO_GRAPH), and the launch accounting is the same."""
    g = golden("1k_bodies")
    res = []
    for flags in (0, nb.TUNE_NO_GRAPH):
        with bh_engine(nb, g["pos_xyzm"], g["vel"], tune_flags=flags) as e:
            out = [e.accelerations()]
            e.step(3)
            out += list(e.gather())          # closes the step: kick2 alone
            e.step(1)
            e.step(2)
            out += list(e.gather())
            out.append(e.accelerations())
            e.upload(g["pos_xyzm"], g["vel"])
            e.step(2)
            out += list(e.gather())
            st = e.stats()
            res.append((out, st.kernel_launches, st.force_launches))
    assert all(np.array_equal(a, b) for a, b in zip(res[0][0], res[1][0]))
    assert res[0][1:] == res[1][1:]


@pytest.mark.parametrize("runs", [2, 3, 5, 8])
@pytest.mark.parametrize("n", [2, 100, 1000, 10_000, 50_000])
def test_bh_slice_sort_and_merge_equals_full_sort(nb, runs, n):
    """Multi-rank engines sort the Morton keys of their own slice and merge the all-gathered runs
    (merge path, ties to the left run) instead of radix-sorting all keys on every rank.  The
    single-rank rehearsal (tune_flags bits 24..27 = number of slices) must reproduce the full
    sort's sequence, hence the same tree and bit-identical results -- including equal keys
    (coincident bodies) and runs that are partly or wholly padding."""
    g = golden("10k_bodies_g1")
    rng = np.random.default_rng(n + runs)
    if n <= 10_000:
        pm, vel = g["pos_xyzm"][:n].copy(), g["vel"][:n].copy()
    else:
        pm = np.concatenate([rng.standard_normal((n, 3)), rng.uniform(0.1, 1.0, (n, 1))], axis=1)
        vel = rng.standard_normal((n, 3))
    if n >= 100:
        pm[n // 2, :3] = pm[3, :3]          # coincident pair: equal keys in different runs
        pm[n - 1, :3] = pm[3, :3]
    res = []
    for flags in (0, runs << nb.TUNE_EMULATE_RUNS_SHIFT, (runs << nb.TUNE_EMULATE_RUNS_SHIFT) | nb.TUNE_NO_GRAPH):
        with bh_engine(nb, pm, vel, tune_flags=flags | nb.TUNE_COUNT_VISITS) as e:
            acc = e.accelerations()
            ti = e.tree_stats()
            e.step(3)
            p, v = e.gather()
        res.append((acc, p, v, ti.nodes, ti.depth, ti.interactions, ti.node_visits))
    for other in res[1:]:
        for a, b in zip(res[0][:3], other[:3]):
            assert np.array_equal(a, b)
        assert res[0][3:] == other[3:]
