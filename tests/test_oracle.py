"""The oracle (oracle/nbody_oracle.c, oracle/bh_oracle.c) pinned against the reference:
 * the reference's only known-answer test (common/get_energy.c:49-59);
 * tests/golden/*.npz, produced by the reference's own compiled sources
   (tests/golden/make_golden.py);
 * bit-for-bit against oracle/_ref (the reference compiled in place), when present.
CPU only.
"""
import hashlib
import os

import numpy as np
import pytest

import oracle as O
from conftest import golden

CASES = ["10_bodies", "20_bodies_g1", "33_bodies", "100_bodies_g1", "500_bodies_g1", "1k_bodies",
         "sun_earth"]
needs_ref = pytest.mark.skipif(not O.have_ref(), reason="oracle/_ref not built (no /root/reference)")


def test_get_energy_known_answer(oracle_lib):
    # inputs of common/get_energy.c:49-59; the reference prints
    # "KE: 1629932.000000000000000, PE: -207.735345789164597"
    vel = np.array([5, 65, 105, -18, 9, -89, -64, 23, 32, 45, 45, 45, -29, -71, -61.0])
    pos = np.array([90, 44, -32, 33, 100, 9, -84, -4, 0, 103, 182, 71, 9, 2, -103.0])
    mass = np.array([103, 50, 88, 34, 59.0])
    import ctypes as C
    ke, pe = C.c_double(), C.c_double()
    oracle_lib.orc_get_energy_flat(pos, vel, mass, 0.7, 5, C.byref(ke), C.byref(pe))
    assert "KE: %.15f, PE: %.15f" % (ke.value, pe.value) == \
        "KE: 1629932.000000000000000, PE: -207.735345789164597"
    kat = str(golden("programs")["get_energy_kat_stdout"]).strip()
    assert kat == "KE: 1629932.000000000000000, PE: -207.735345789164597"


@pytest.mark.parametrize("name", CASES + ["10k_bodies_g1"])
def test_acceleration_bitwise_vs_golden(oracle_lib, name):
    g = golden(name)
    acc = O.acceleration(oracle_lib, g["pos_xyzm"], "orc_acceleration_mt", 8)
    assert np.array_equal(acc, g["acc0"])


@pytest.mark.parametrize("name", CASES)
def test_propagation_bitwise_vs_golden(oracle_lib, name):
    g = golden(name)
    for key in g.files:
        if not key.startswith("pos_") or key == "pos_xyzm":
            continue
        k = int(key[4:])
        pm, v = O.propagate(oracle_lib, g["pos_xyzm"], g["vel"], k, g["dt"], threads=8)
        assert np.array_equal(pm, g["pos_%d" % k]) and np.array_equal(v, g["vel_%d" % k])
        ke, pe = O.get_energy(oracle_lib, pm, v)
        assert ke == float(g["ke_%d" % k]) and pe == float(g["pe_%d" % k])


def test_serial_and_threaded_oracle_agree(oracle_lib):
    g = golden("100_bodies_g1")
    a, b = O.propagate(oracle_lib, g["pos_xyzm"], g["vel"], 10, g["dt"], threads=1)
    c, d = O.propagate(oracle_lib, g["pos_xyzm"], g["vel"], 10, g["dt"], threads=4)
    assert np.array_equal(a, c) and np.array_equal(b, d)


def test_frame_file_md5_of_headline_cpu_config(oracle_lib, tmp_path):
    """BASELINE.json configs[0]: 1k bodies, t = 0..1, dt = 0.01 -> 101 000 lines whose md5 the
    reference's exh_serial and exh_mp both produce."""
    g, prog = golden("1k_bodies"), golden("programs")
    out = str(tmp_path / "frames.csv")
    O.propagate(oracle_lib, g["pos_xyzm"], g["vel"], 100, g["dt"], frames=out, threads=8)
    data = open(out, "rb").read()
    assert data.count(b"\n") == int(prog["exh_serial_results_lines"]) == 101000
    assert hashlib.md5(data).hexdigest() == str(prog["exh_serial_results_md5"])
    assert str(prog["exh_serial_results_md5"]) == str(prog["exh_mp_results_md5"])
    assert "time steps: 100, ents: 1000" in str(prog["exh_serial_results_stdout"])


def test_parse_times_float_semantics(oracle_lib):
    import ctypes as C
    s, e, d = C.c_float(), C.c_float(), C.c_float()
    n = oracle_lib.orc_parse_times(b"0", b"1", b"0.01", C.byref(s), C.byref(e), C.byref(d))
    assert n == 100 and d.value == np.float32(0.01)
    assert float(d.value) != 0.01  # the effective step is (double)0.01f
    assert oracle_lib.orc_parse_times(b"0", b"0.05", b"0.01", C.byref(s), C.byref(e), C.byref(d)) == 5


@pytest.mark.parametrize("name", ["10_bodies", "100_bodies_g1", "1k_bodies", "10k_bodies_g1"])
def test_barnes_hut_bitwise_vs_golden(oracle_lib, name):
    g = golden(name)
    n = g["pos_xyzm"].shape[0]
    acc = np.zeros((n, 3))
    st = np.zeros(5)
    oracle_lib.orc_bh_acceleration(n, np.ascontiguousarray(g["pos_xyzm"]), acc, st.ctypes.data)
    assert np.array_equal(acc, g["bh_acc0"])
    assert st[0] == g["bh_stats"][0] and st[1] == g["bh_stats"][1] and st[2] == g["bh_stats"][2]


def test_barnes_hut_error_envelope(oracle_lib):
    """THETA = 0.5 against the direct sum on identical positions (SURVEY 8c envelope)."""
    g = golden("1k_bodies")
    from conftest import rel_err
    err = rel_err(g["bh_acc0"], g["acc0"])
    assert np.median(err) <= 4e-3 and np.percentile(err, 99) <= 2e-2 and err.max() <= 1.2e-1


def test_f32_model_close_to_fp64(oracle_lib):
    g = golden("1k_bodies")
    from conftest import rel_err
    pm32 = g["pos_xyzm"].astype(np.float32).astype(np.float64)
    a32 = O.acceleration(oracle_lib, pm32, "orc_acceleration_f32")
    a64 = O.acceleration(oracle_lib, pm32)
    assert rel_err(a32, a64).max() <= 1e-5


# ------------------------------------------------ against the reference ----
@needs_ref
@pytest.mark.parametrize("name", ["33_bodies", "1k_bodies"])
def test_bitwise_vs_compiled_reference(oracle_lib, name):
    g = golden(name)
    rs = O.load_ref("exh_serial")
    pm = np.ascontiguousarray(g["pos_xyzm"])
    assert np.array_equal(O.acceleration(oracle_lib, pm), O.acceleration(rs, pm, "ref_acceleration"))
    a, b = O.propagate(oracle_lib, pm, g["vel"], 10, g["dt"])
    p2, v2 = pm.copy(), np.array(g["vel"], order="C")
    rs.ref_propagate(pm.shape[0], p2, v2, 10, np.float32(g["dt"]), b"/dev/null")
    assert np.array_equal(a, p2) and np.array_equal(b, v2)
    rb = O.load_ref("bh_serial")
    bh_ref, bh = np.zeros((pm.shape[0], 3)), np.zeros((pm.shape[0], 3))
    rb.ref_bh_acceleration(pm.shape[0], pm, bh_ref, None)
    oracle_lib.orc_bh_acceleration(pm.shape[0], pm, bh, None)
    assert np.array_equal(bh, bh_ref)


@needs_ref
def test_reader_matches_reference_reader(oracle_lib, tmp_path):
    g = golden("sun_earth")
    path = str(tmp_path / "sun_earth.csv")
    open(path, "wb").write(bytes(g["csv_text"]))  # spaces after commas, no trailing newline
    rc, p, v = O.read_bodies(oracle_lib, path)
    rs = O.load_ref("exh_serial")
    _, p2, v2 = O.read_bodies(rs, path, "ref_read_bodies")
    assert rc == 0 and np.array_equal(p, p2) and np.array_equal(v, v2)
    assert np.array_equal(p, g["pos_xyzm"]) and p.shape == (9, 4)
