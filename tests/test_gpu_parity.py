"""GPU parity tests: the CUDA engine, called through its C ABI, against the
reference's own numbers (tests/golden/*.npz, produced by the reference's
compiled sources) and against the oracle on seeded inputs.

Tolerances (written here as the test's contract):
  FP64  per-body relative error (|d| / |oracle vector|) of positions/velocities
        <= 1e-12 after K <= 10 steps, <= 1e-10 after K = 100 steps;
        accelerations <= 1e-12; KE / PE <= 1e-9 relative.
  FP32  accelerations <= 1e-5 relative to the FP64 reference on the same
        (float-rounded) positions; state <= 1e-5 for K <= 10; KE / PE <= 1e-5.
The engine differs from the reference only by FMA contraction, summation
order and rsqrt^3 in place of pow(x,-1.5); the reference itself moves by
2.7e-11 under -ffp-contract=fast over 100 steps of the 1k fixture.
"""
import numpy as np
import pytest

from conftest import golden, rel_err

pytestmark = pytest.mark.gpu

FP64_CASES = ["10_bodies", "20_bodies_g1", "33_bodies", "100_bodies_g1", "500_bodies_g1",
              "1k_bodies", "sun_earth"]


def tol64(k):
    return 1e-12 if k <= 10 else 1e-10


@pytest.mark.parametrize("name", FP64_CASES + ["10k_bodies_g1"])
def test_fp64_accelerations_match_reference(nb, name):
    g = golden(name)
    with nb.Engine(g["pos_xyzm"], g["vel"], precision=nb.PREC_FP64) as e:
        acc = e.accelerations()
    assert rel_err(acc, g["acc0"]).max() <= 1e-12


@pytest.mark.parametrize("name", FP64_CASES)
def test_fp64_state_after_k_steps(nb, name):
    g = golden(name)
    ks = sorted(int(k[4:]) for k in g.files if k.startswith("pos_") and k != "pos_xyzm")
    for k in ks:
        with nb.Engine(g["pos_xyzm"], g["vel"], precision=nb.PREC_FP64, dt=g["dt"]) as e:
            e.step(k)
            pos, vel = e.gather()
            st = e.stats()
        assert st.force_launches == k + 1  # the reference's n_steps + 1 force evaluations
        assert rel_err(pos, g["pos_%d" % k]).max() <= tol64(k), (name, k)
        assert rel_err(vel, g["vel_%d" % k]).max() <= tol64(k), (name, k)
        assert np.array_equal(pos[:, 3], g["pos_xyzm"][:, 3])  # masses ride along unchanged


def test_fp64_10k_state_after_10_steps(nb):
    g = golden("10k_bodies_g1")
    with nb.Engine(g["pos_xyzm"], g["vel"], precision=nb.PREC_FP64, dt=g["dt"]) as e:
        e.step(10)
        pos, vel = e.gather()
    # 10k bodies of mass 0.2 is a violent system: the reference's own serial code moves by
    # 9.9e-13 (max) / 7e-16 (median) over these 10 steps when gcc contracts FMAs
    # (-O3 -march=x86-64-v3), so the bound here is 1e-11 max with a 1e-14 median.
    ep, ev = rel_err(pos, g["pos_10"]), rel_err(vel, g["vel_10"])
    assert ep.max() <= 1e-11 and ev.max() <= 1e-11
    assert np.median(ep) <= 1e-14 and np.median(ev) <= 1e-14


@pytest.mark.parametrize("name", ["100_bodies_g1", "1k_bodies"])
def test_fp64_energy_matches_get_energy(nb, name):
    g = golden(name)
    with nb.Engine(g["pos_xyzm"], g["vel"], precision=nb.PREC_FP64, dt=g["dt"]) as e:
        ke0, pe0 = e.energy()
        e.step(100)
        ke1, pe1 = e.energy()
    assert abs(ke0 - g["ke_0"]) <= 1e-12 * abs(g["ke_0"])
    assert abs(pe0 - g["pe_0"]) <= 1e-12 * abs(g["pe_0"])
    assert abs(ke1 - g["ke_100"]) <= 1e-9 * abs(g["ke_100"])
    assert abs(pe1 - g["pe_100"]) <= 1e-9 * abs(g["pe_100"])


@pytest.mark.parametrize("n", [1, 2, 255, 256, 257, 700, 1300, 5000])
@pytest.mark.parametrize("prec", ["fp64", "fp32"])
def test_energy_kernel_tile_pairing(nb, oracle_lib, n, prec):
    """The energy kernel pairs row tile b with tile (last - b) per CTA: odd and even tile counts,
    the middle tile done once, ragged last tiles, coincident bodies (r = 0 contributes nothing,
    common/get_energy.c:33-36)."""
    import oracle as O
    rng = np.random.default_rng(n)
    pm = np.concatenate([rng.standard_normal((n, 3)), rng.uniform(0.5, 2, (n, 1))], axis=1)
    vel = rng.standard_normal((n, 3))
    if n >= 256:
        pm[200, :3] = pm[3, :3]           # a coincident pair inside one tile
        pm[n - 1, :3] = pm[0, :3]         # and one across the first and the last tile
    if prec == "fp32":
        pm, vel = pm.astype(np.float32).astype(np.float64), vel.astype(np.float32).astype(np.float64)
    P = nb.PREC_FP64 if prec == "fp64" else nb.PREC_FP32
    with nb.Engine(pm, vel, precision=P, G=0.7) as e:
        ke, pe = e.energy()
    oke, ope = O.get_energy(oracle_lib, pm, vel, G=0.7)
    assert abs(ke - oke) <= 1e-13 * abs(oke)
    assert abs(pe - ope) <= 1e-12 * max(abs(ope), 1e-300)


def test_get_energy_known_answer(nb):
    # common/get_energy.c:49-59, the reference's only known-answer test (G = 0.7)
    vel = np.array([5, 65, 105, -18, 9, -89, -64, 23, 32, 45, 45, 45, -29, -71, -61.0]).reshape(5, 3)
    pos = np.array([90, 44, -32, 33, 100, 9, -84, -4, 0, 103, 182, 71, 9, 2, -103.0]).reshape(5, 3)
    mass = np.array([103, 50, 88, 34, 59.0])
    pm = np.concatenate([pos, mass[:, None]], axis=1)
    with nb.Engine(pm, vel, G=0.7) as e:
        ke, pe = e.energy()
    assert ke == 1629932.0
    assert abs(pe - (-207.735345789164597)) <= 1e-13 * 207.7


def test_stepping_patterns_agree(nb):
    """step(6)+gather == 6 x (step(1), gather with velocities): the stored-acceleration
    kick+drift path and the fused path are the same arithmetic."""
    g = golden("100_bodies_g1")
    with nb.Engine(g["pos_xyzm"], g["vel"], dt=g["dt"]) as a:
        a.step(6)
        pa, va = a.gather()
        la = a.stats().force_launches
    with nb.Engine(g["pos_xyzm"], g["vel"], dt=g["dt"]) as b:
        for _ in range(6):
            b.step(1)
            pb, vb = b.gather()
        lb = b.stats().force_launches
    assert np.array_equal(pa, pb) and np.array_equal(va, vb)
    assert la == 7 and lb == 7
    with nb.Engine(g["pos_xyzm"], g["vel"], dt=g["dt"]) as c:  # positions-only gathers cost nothing
        for _ in range(6):
            c.step(1)
            c.gather(with_velocities=False)
        pc, vc = c.gather()
        assert c.stats().force_launches == 7
    assert np.array_equal(pa, pc) and np.array_equal(va, vc)


def test_upload_resets_state(nb):
    g = golden("33_bodies")
    with nb.Engine(g["pos_xyzm"], g["vel"], dt=g["dt"]) as e:
        e.step(10)
        p1, v1 = e.gather()
        e.upload(g["pos_xyzm"], g["vel"])
        e.step(10)
        p2, v2 = e.gather()
    assert np.array_equal(p1, p2) and np.array_equal(v1, v2)
    assert rel_err(p1, g["pos_10"]).max() <= 1e-12


@pytest.mark.parametrize("n", [1, 2, 3, 5, 31, 33, 127, 129, 255, 257])
@pytest.mark.parametrize("prec", ["fp64", "fp32"])
def test_ragged_sizes(nb, oracle_lib, n, prec):
    """tile / quad / warp boundaries: padding bodies must contribute exactly nothing."""
    import oracle as O
    rng = np.random.default_rng(100 + n)
    pm = np.concatenate([rng.standard_normal((n, 3)), rng.uniform(0.1, 1.0, (n, 1))], axis=1)
    vel = rng.standard_normal((n, 3))
    P = nb.PREC_FP64 if prec == "fp64" else nb.PREC_FP32
    if prec == "fp32":
        pm = pm.astype(np.float32).astype(np.float64)
        vel = vel.astype(np.float32).astype(np.float64)
    ref_acc = O.acceleration(oracle_lib, pm)
    rp, rv = O.propagate(oracle_lib, pm, vel, 3, 0.01)
    with nb.Engine(pm, vel, precision=P) as e:
        acc = e.accelerations()
        e.step(3)
        pos, v = e.gather()
    tol = 1e-12 if prec == "fp64" else 1e-5
    scale = np.abs(ref_acc).max() if n > 1 else 1.0
    assert np.abs(acc - ref_acc).max() <= tol * max(scale, 1e-30)
    assert rel_err(pos, rp).max() <= tol and rel_err(v, rv).max() <= tol


def test_empty_input(nb):
    with nb.Engine(np.zeros((0, 4)), np.zeros((0, 3))) as e:
        e.step(2)
        pos, vel = e.gather()
    assert pos.shape == (0, 4) and vel.shape == (0, 3)


# ------------------------------------------------------------------ FP32 ---
@pytest.mark.parametrize("name", ["100_bodies_g1", "1k_bodies", "10k_bodies_g1"])
def test_fp32_accelerations(nb, name):
    """FP32 engine vs the FP64 reference acceleration on the float-rounded positions."""
    import oracle as O
    g = golden(name)
    pm32 = g["pos_xyzm"].astype(np.float32).astype(np.float64)
    ref = O.acceleration(O.load_oracle(), pm32, "orc_acceleration_mt", 8)
    with nb.Engine(pm32, g["vel"], precision=nb.PREC_FP32) as e:
        acc = e.accelerations()
    err = rel_err(acc, ref)
    assert err.max() <= 1e-5, err.max()
    assert np.median(err) <= 1e-6


@pytest.mark.parametrize("name", ["100_bodies_g1", "1k_bodies"])
def test_fp32_state_after_10_steps(nb, name):
    import oracle as O
    g = golden(name)
    pm32 = g["pos_xyzm"].astype(np.float32).astype(np.float64)
    v32 = g["vel"].astype(np.float32).astype(np.float64)
    rp, rv = O.propagate(O.load_oracle(), pm32, v32, 10, g["dt"], threads=8)
    with nb.Engine(pm32, v32, precision=nb.PREC_FP32, dt=g["dt"]) as e:
        e.step(10)
        pos, vel = e.gather()
        ke, pe = e.energy()
    assert rel_err(pos, rp).max() <= 1e-5
    assert rel_err(vel, rv).max() <= 1e-5
    rke, rpe = O.get_energy(O.load_oracle(), rp, rv)
    assert abs(ke - rke) <= 1e-5 * abs(rke) and abs(pe - rpe) <= 1e-5 * abs(rpe)


@pytest.mark.parametrize("prec", ["fp64", "fp32"])
def test_split_and_unsplit_work_lists_agree(nb, prec):
    """The evenly cut work list (partial j-ranges + last-arriver sum) against one
    CTA per i-block over all j, and the packed f32x2 path against scalar FP32."""
    g = golden("10k_bodies_g1")
    P = nb.PREC_FP64 if prec == "fp64" else nb.PREC_FP32
    outs = []
    for flags in ([0, nb.TUNE_NO_SPLIT] + ([nb.TUNE_SCALAR_FP32] if prec == "fp32" else [])):
        with nb.Engine(g["pos_xyzm"], g["vel"], precision=P, tune_flags=flags) as e:
            outs.append(e.accelerations())
    tol = 1e-13 if prec == "fp64" else 2e-6
    for o in outs[1:]:
        assert rel_err(o, outs[0]).max() <= tol


@pytest.mark.parametrize("variant", [1, 2, 3])
@pytest.mark.parametrize("prec", ["fp64", "fp32"])
def test_all_size_classes(nb, variant, prec):
    """every kernel size class on the same input"""
    g = golden("1k_bodies")
    P = nb.PREC_FP64 if prec == "fp64" else nb.PREC_FP32
    with nb.Engine(g["pos_xyzm"], g["vel"], precision=P, tune_flags=variant << 8) as e:
        acc = e.accelerations()
    assert rel_err(acc, g["acc0"]).max() <= (1e-12 if prec == "fp64" else 1e-5)


@pytest.mark.parametrize("variant", [4, 5, 6, 7, 9, 10, 12])
@pytest.mark.parametrize("masses", ["equal", "spread"])
@pytest.mark.parametrize("n", [5, 1000, 10_000, 33_000])
def test_fp32_ipacked_size_classes(nb, oracle_lib, variant, masses, n):
    """The i-packed FP32 kernels (8 i-bodies per thread; 1x16, 1x12, 2x8, 1x8 warps per SM), forced
    through tune_flags on inputs from a few bodies (one ragged i-block) to several i-blocks whose
    j-ranges are cut between CTAs: accelerations against the oracle, 3 leapfrog steps against the
    j-packed kernel (variant 1), equal and general masses."""
    import oracle as O
    g = golden("10k_bodies_g1")
    rng = np.random.default_rng(n)
    if n <= 10_000:
        pm, vel = g["pos_xyzm"][:n].copy(), g["vel"][:n].copy()
    else:
        pm = np.concatenate([rng.standard_normal((n, 3)), np.full((n, 1), 0.2)], axis=1)
        vel = rng.standard_normal((n, 3))
    if masses == "spread":
        pm[:, 3] = rng.uniform(0.1, 0.3, n)
    pm = pm.astype(np.float32).astype(np.float64)
    vel = vel.astype(np.float32).astype(np.float64)
    with nb.Engine(pm, vel, precision=nb.PREC_FP32, tune_flags=variant << 8) as e:
        acc = e.accelerations()
        e.step(3)
        p, v = e.gather()
        assert e.stats().uniform_mass == (1 if masses == "equal" else 0)
    with nb.Engine(pm, vel, precision=nb.PREC_FP32, tune_flags=1 << 8) as e:
        acc1 = e.accelerations()
        e.step(3)
        p1, v1 = e.gather()
    if n <= 10_000:
        ref = O.acceleration(oracle_lib, pm, "orc_acceleration_mt", 8)
        assert rel_err(acc, ref).max() <= 1e-5
    assert rel_err(acc, acc1).max() <= 2e-6
    assert rel_err(p, p1).max() <= 2e-5 and rel_err(v, v1).max() <= 2e-5
    assert np.array_equal(p[:, 3], pm[:, 3])


@pytest.mark.parametrize("variant", [1, 2, 3, 4, 5, 6, 7, 8])
@pytest.mark.parametrize("masses", ["equal", "spread"])
@pytest.mark.parametrize("n", [3, 1000, 2500, 10_000])
def test_fp64_size_classes_and_equal_masses(nb, oracle_lib, variant, masses, n):
    """Every FP64 direct kernel (2 and 4 i-bodies per thread; the equal-mass form that drops m_j
    from the loop and parks the padding at 1e150, and the general form), forced through
    tune_flags: accelerations against the oracle at 1e-12, 10 leapfrog steps against the reference's
    propagation at 1e-12, masses returned untouched."""
    import oracle as O
    g = golden("10k_bodies_g1")
    pm, vel = g["pos_xyzm"][:n].copy(), g["vel"][:n].copy()
    if masses == "spread":
        pm[:, 3] = np.random.default_rng(n).uniform(0.1, 0.3, n)
    with nb.Engine(pm, vel, precision=nb.PREC_FP64, tune_flags=variant << 8) as e:
        st = e.stats()
        assert st.kernel_variant == variant - 1
        acc = e.accelerations()
        assert e.stats().uniform_mass == (1 if masses == "equal" else 0)
        e.step(10)
        p, v = e.gather()
    ref = O.acceleration(oracle_lib, pm, "orc_acceleration_mt", 8)
    assert rel_err(acc, ref).max() <= 1e-12
    if n <= 2500:
        rp, rv = O.propagate(oracle_lib, pm, vel, 10, 0.01, threads=8)
        assert rel_err(p[:, :3], rp[:, :3]).max() <= 1e-12
        assert rel_err(v, rv).max() <= 1e-12
    assert np.array_equal(p[:, 3], pm[:, 3])
    # the equal-mass form against the general one on the same input
    with nb.Engine(pm, vel, precision=nb.PREC_FP64, tune_flags=(variant << 8) | nb.TUNE_NO_UNIFORM) as e:
        assert e.stats().uniform_mass == 0
        assert rel_err(e.accelerations(), acc).max() <= 1e-13


# -------------------------------------------- size-independent properties ---
@pytest.mark.parametrize("prec,n,masses", [("fp32", 200_000, "spread"), ("fp64", 60_000, "spread"),
                                           ("fp32", 1_000_000, "equal"),    # BASELINE config 3 (bench)
                                           ("fp32", 1_000_000, "spread"),
                                           ("fp64", 100_000, "equal")])     # BASELINE config 2
def test_large_n_properties(nb, prec, n, masses):
    """At sizes the CPU oracle cannot reach in seconds, up to BASELINE.json's full sizes:
    (1) Newton's third law -- sum_i m_i a_i vanishes to rounding; (2) scaling every mass by 2
    doubles every acceleration bit-exactly; (3) a random sample of rows against the oracle."""
    import oracle as O
    rng = np.random.default_rng(17)
    m = rng.uniform(0.5, 1.5, (n, 1)) / n if masses == "spread" else np.full((n, 1), 1.0 / n)
    pm = np.concatenate([rng.standard_normal((n, 3)), m], axis=1)
    if prec == "fp32":
        pm = pm.astype(np.float32).astype(np.float64)
    P = nb.PREC_FP64 if prec == "fp64" else nb.PREC_FP32
    vel = np.zeros((n, 3))
    with nb.Engine(pm, vel, precision=P) as e:
        acc = e.accelerations()
    pm2 = pm.copy()
    pm2[:, 3] *= 2
    with nb.Engine(pm2, vel, precision=P) as e:
        acc2 = e.accelerations()
    assert np.array_equal(acc2, 2 * acc)
    net = (pm[:, 3:4] * acc).sum(axis=0)
    typical = (pm[:, 3] * np.linalg.norm(acc, axis=1)).sum()
    assert np.abs(net).max() <= (1e-12 if prec == "fp64" else 2e-6) * typical
    rows = rng.choice(n, 64, replace=False)
    lib = O.load_oracle()
    for r in rows[:16]:
        ref = np.zeros((1, 3))
        lib.orc_acceleration_slice(n, pm, int(r), int(r) + 1, ref)
        err = np.linalg.norm(acc[r] - ref[0]) / np.linalg.norm(ref[0])
        assert err <= (1e-12 if prec == "fp64" else 1e-5)


@pytest.mark.parametrize("n", [7, 1000, 10_000])
def test_fp32_uniform_mass_path(nb, n):
    """Equal masses: the FP32 kernel drops m_j from the inner loop (padding parked far away
    instead of massless).  Same answers as the general kernel, to FP32 rounding."""
    import oracle as O
    g = golden("10k_bodies_g1")           # mass 0.2 everywhere
    pm = g["pos_xyzm"][:n].astype(np.float32).astype(np.float64)
    vel = g["vel"][:n].astype(np.float32).astype(np.float64)
    ref = O.acceleration(O.load_oracle(), pm, "orc_acceleration_mt", 8)
    outs = []
    for flags in (0, nb.TUNE_NO_UNIFORM):
        with nb.Engine(pm, vel, precision=nb.PREC_FP32, tune_flags=flags) as e:
            acc = e.accelerations()
            e.step(3)
            outs.append((acc,) + e.gather())
    for acc, p, v in outs:
        assert rel_err(acc, ref).max() <= 1e-5
        assert np.isfinite(p).all() and np.isfinite(v).all()
        assert np.array_equal(p[:, 3], pm[:, 3])
    assert rel_err(outs[0][0], outs[1][0]).max() <= 2e-6
    assert rel_err(outs[0][1], outs[1][1]).max() <= 2e-6


@pytest.mark.parametrize("prec", ["fp64", "fp32"])
def test_frame_pipeline_equals_blocking_gather(nb, prec):
    """nb_frame_begin / nb_frame_wait (asynchronous read-back, two snapshots in flight while
    later steps run) deliver exactly the frames the blocking nb_gather loop delivers."""
    g = golden("500_bodies_g1")
    precision = nb.PREC_FP32 if prec == "fp32" else nb.PREC_FP64
    want = []
    with nb.Engine(g["pos_xyzm"], g["vel"], precision=precision, dt=0.01) as e:
        for _ in range(7):
            e.step(1)
            want.append(e.gather(with_velocities=False)[0])
        pos_a, vel_a = e.gather()
    with nb.Engine(g["pos_xyzm"], g["vel"], precision=precision, dt=0.01) as e:
        buf = [e.frame_buffer(), e.frame_buffer()]
        got = []
        for t in range(8):
            if t < 7:
                e.step(1)
                e.frame_begin(buf[t & 1])
            if t > 0:
                e.frame_wait(buf[(t - 1) & 1])
                got.append(buf[(t - 1) & 1].copy())
        e.frame_wait()
        pos_b, vel_b = e.gather()
        launches = e.stats().force_launches
    assert len(got) == 7 and all(np.array_equal(a, b) for a, b in zip(want, got))
    assert np.array_equal(pos_a, pos_b) and np.array_equal(vel_a, vel_b)
    assert launches == 8      # the pipeline adds no force evaluations: n + 1 for n steps
    # pageable destination: still correct (the driver stages the copy)
    with nb.Engine(g["pos_xyzm"], g["vel"], precision=precision, dt=0.01) as e:
        out = np.empty((500, 4))
        e.step(1)
        e.frame_begin(out)
        e.step(1)
        e.frame_wait(out)
        assert np.array_equal(out, want[0])


def test_direct_sum_refuses_zero_softening(nb):
    """The direct kernels include j = i like the reference (serial/exhaustive.c:163) and rely on
    eps2 > 0 to make that term exactly 0; eps2 = 0 would give 0 * inf = NaN for every body."""
    g = golden("10_bodies")
    for prec in (nb.PREC_FP64, nb.PREC_FP32):
        with pytest.raises(nb.NbodyError) as ei:
            nb.Engine(g["pos_xyzm"], g["vel"], precision=prec, eps2=0.0)
        assert ei.value.code == -1
    # Barnes-Hut skips the body's own leaf by index (:411) and works without softening
    with nb.Engine(g["pos_xyzm"], g["vel"], precision=nb.PREC_FP64, algorithm=nb.ALGO_BARNES_HUT,
                   eps2=0.0) as e:
        assert np.isfinite(e.accelerations()).all()


@pytest.mark.parametrize("prec", ["fp64", "fp32"])
def test_slice_io_on_one_rank_is_the_full_io(nb, prec):
    """world = 1: nb_upload_slice / nb_gather_slice are nb_upload / nb_gather."""
    g = golden("1k_bodies")
    P = nb.PREC_FP64 if prec == "fp64" else nb.PREC_FP32
    with nb.Engine(g["pos_xyzm"], g["vel"], precision=P, dt=g["dt"]) as e:
        assert e.io_slice() == (0, 1000)
        e.step(2)
        p1, v1 = e.gather()
        e.upload_slice(g["pos_xyzm"], g["vel"])
        e.step(2)
        p2, v2 = e.gather_slice()
    assert np.array_equal(p1, p2) and np.array_equal(v1, v2)


def test_equal_mass_detection_runs_on_the_device(nb):
    """The mass class (equal masses -> m_j hoisted out of the FP32 loop) is decided by a kernel on
    the staged rows; stats().uniform_mass reports it, and a re-upload with other masses flips it."""
    rng = np.random.default_rng(5)
    n = 777
    pm = np.concatenate([rng.standard_normal((n, 3)), np.full((n, 1), 0.25)], axis=1)
    vel = np.zeros((n, 3))
    with nb.Engine(pm, vel, precision=nb.PREC_FP32) as e:
        assert e.stats().uniform_mass == 1
        a_uni = e.accelerations()
        pm2 = pm.copy()
        pm2[n - 1, 3] = 0.5
        e.upload(pm2, vel)
        assert e.stats().uniform_mass == 0
        e.upload(pm, vel)
        assert e.stats().uniform_mass == 1
        assert np.array_equal(e.accelerations(), a_uni)
    with nb.Engine(pm, vel, precision=nb.PREC_FP32, tune_flags=nb.TUNE_NO_UNIFORM) as e:
        assert e.stats().uniform_mass == 0
        assert rel_err(e.accelerations(), a_uni).max() <= 2e-6
    with nb.Engine(pm, vel, precision=nb.PREC_FP64) as e:
        assert e.stats().uniform_mass == 1
        a64 = e.accelerations()
        pm2[n - 1, 3] = 0.25 * (1 + 2.0 ** -40)   # equal as floats, not as doubles
        e.upload(pm2, vel)
        assert e.stats().uniform_mass == 0
        assert rel_err(e.accelerations(), a64).max() <= 1e-11
    with nb.Engine(pm, vel, precision=nb.PREC_FP64, algorithm=nb.ALGO_BARNES_HUT) as e:
        assert e.stats().uniform_mass == 0            # the tree needs the masses: never hoisted
