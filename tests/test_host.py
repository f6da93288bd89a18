"""CPU-side tests of the drop-in boundary: the C ABI library loads and exports every
symbol include/nbody_engine.h declares (no compute without a GPU), it refuses to run
without a device, and the C host I/O (CSV reader, argv times, frame writer) behaves like
the reference's (serial/exhaustive.c:43-53,79-146,208-238).  No GPU needed.
"""
import ctypes as C
import os
import re

import numpy as np
import pytest

import oracle as O
from conftest import ROOT, golden


def test_abi_exports_every_declared_symbol(nb):
    hdr = open(os.path.join(ROOT, "include", "nbody_engine.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = sorted(set(re.findall(r"\b(nb_[a-z_]+)\s*\(", hdr)))
    assert len(declared) >= 15
    lib = nb.load_library()
    for sym in declared:
        assert getattr(lib, sym) is not None, sym
    assert sorted(nb.ABI_SYMBOLS) == declared
    assert nb.load_library().nb_version().decode().startswith("nbody_b200")


def test_no_cpu_fallback(nb):
    """Without a CUDA device the engine must fail loudly, not compute on the host."""
    if nb.device_count() > 0:
        pytest.skip("a GPU is present")
    g = golden("10_bodies")
    with pytest.raises(nb.NbodyError) as ei:
        nb.Engine(g["pos_xyzm"], g["vel"])
    assert ei.value.code == -5  # NB_E_NODEVICE
    assert "no CPU fallback" in str(ei.value)


def test_config_defaults_are_the_reference_constants(nb):
    cfg = nb.NbConfig()
    nb.load_library().nb_config_default(C.byref(cfg))
    assert (cfg.precision, cfg.algorithm, cfg.world, cfg.rank) == (nb.PREC_FP64, nb.ALGO_DIRECT, 1, 0)
    assert cfg.eps2 == 0.01 and cfg.G == 1.0 and cfg.theta == 0.5   # serial/exhaustive.c:24,172
    assert cfg.dt == float(np.float32(0.01))                        # dt is a float: :43,51


def test_product_package_does_not_touch_the_oracle():
    pkg = os.path.join(ROOT, "n-body-simulation_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".c", ".h", ".cu", ".cuh")):
                src = open(os.path.join(dp, f)).read()
                assert "oracle" not in src, os.path.join(dp, f)


# ------------------------------------------------------------------ reader ---
def test_reader_accepts_what_the_reference_accepts(nb, tmp_path):
    g = golden("sun_earth")
    path = str(tmp_path / "sun_earth.csv")
    open(path, "wb").write(bytes(g["csv_text"]))     # spaces after commas, no trailing newline
    pos, vel = nb.read_bodies(path)
    assert np.array_equal(pos, g["pos_xyzm"]) and np.array_equal(vel, g["vel"])
    open(path, "wb").write(bytes(g["csv_text"]) + b"\n")
    pos2, _ = nb.read_bodies(path)
    assert np.array_equal(pos2, pos)


def test_reader_matches_oracle_reader_on_all_fixtures(nb, oracle_lib, tmp_path):
    for name in ["10_bodies", "100_bodies_g1", "1k_bodies"]:
        g = golden(name)
        rows = np.concatenate([g["pos_xyzm"][:, :3], g["vel"], g["pos_xyzm"][:, 3:4]], axis=1)
        path = str(tmp_path / (name + ".csv"))
        open(path, "w").write("".join(",".join(repr(float(x)) for x in r) + "\n" for r in rows))
        pos, vel = nb.read_bodies(path)
        rc, opos, ovel = O.read_bodies(oracle_lib, path)
        assert rc == 0 and np.array_equal(pos, opos) and np.array_equal(vel, ovel)
        assert np.array_equal(pos, g["pos_xyzm"])


def test_reader_errors(nb, tmp_path):
    with pytest.raises(OSError):
        nb.read_bodies(str(tmp_path / "missing.csv"))
    bad = tmp_path / "bad.csv"
    for text in ["1,2,3,4,5,6\n", "1,2,3,4,5,6,7\nx\n", "1,2,3,4,5,6,7,8\n", "1;2;3;4;5;6;7\n"]:
        bad.write_text(text)
        with pytest.raises(ValueError):
            nb.read_bodies(str(bad))
    empty = tmp_path / "empty.csv"
    empty.write_text("")
    pos, vel = nb.read_bodies(str(empty))
    assert pos.shape == (0, 4) and vel.shape == (0, 3)


def test_parse_times_is_float_arithmetic(nb, oracle_lib):
    """n_steps = (end - start) / dt evaluated in float and truncated (serial/exhaustive.c:49-53)"""
    for args in [("0", "1", "0.01"), ("0", "0.1", "0.01"), ("0", "0.3", "0.1"), ("1", "3.5", "0.25"),
                 ("0", "10", "0.001"), ("0.5", "0.5", "0.01")]:
        s, e, d, n = nb.parse_times(*args)
        fs, fe, fd = C.c_float(), C.c_float(), C.c_float()
        on = oracle_lib.orc_parse_times(*(a.encode() for a in args), C.byref(fs), C.byref(fe), C.byref(fd))
        assert (s, e, d, n) == (fs.value, fe.value, fd.value, on)
        assert n == int(np.float32(np.float32(e) - np.float32(s)) / np.float32(d))
    assert nb.parse_times("0", "1", "0.01")[3] == 100


# ------------------------------------------------------------- frame writer ---
def test_format_lf_equals_printf():
    from conftest import load_package
    h = load_package().load_host_library()
    rng = np.random.default_rng(5)
    vals = np.concatenate([
        rng.standard_normal(4000), rng.standard_normal(2000) * 1e6, rng.standard_normal(2000) * 1e-6,
        rng.standard_normal(500) * 1e15, np.array([0.0, -0.0, 0.5e-6, 1.5e-6, 2.5e-6, -0.5e-6,
                                                    0.9999995, 0.9999994999, 999999.9999995, 1e22,
                                                    123456789.123456789, -4.5e12, 2e30, 1e-300]),
        np.round(rng.standard_normal(2000), 6) + 0.5e-6])
    buf = C.create_string_buffer(512)
    for v in vals:
        k = h.nbh_format_lf(buf, float(v))
        assert buf.raw[:k].decode() == "%f" % v, v


def test_format_frame_is_the_reference_line_format(nb):
    g = golden("programs")
    want = str(g["frames_100_bodies_g1_5steps"]).splitlines()[:100]
    got = nb.format_frame(golden("100_bodies_g1")["pos_xyzm"]).decode().splitlines()
    assert got == want       # "%lu,%lf,%lf,%lf,%lf\n" of frame 0 (serial/exhaustive.c:212-214)


def test_write_bodies_round_trips(nb, tmp_path):
    g = golden("33_bodies")
    h = nb.load_host_library()
    path = str(tmp_path / "ck.csv")
    assert h.nbh_write_bodies(path.encode(), 33, np.ascontiguousarray(g["pos_xyzm"]),
                              np.ascontiguousarray(g["vel"])) == 0
    pos, vel = nb.read_bodies(path)
    assert np.array_equal(pos, g["pos_xyzm"]) and np.array_equal(vel, g["vel"])


def test_slice_bounds_partition(nb):
    for prec in (nb.PREC_FP32, nb.PREC_FP64):
        for n in (0, 1, 5, 1000, 10_000, 1_000_000, 16_000_001):
            for world in (1, 2, 3, 4, 8):
                b = [nb.slice_bounds(n, world, r, prec) for r in range(world)]
                assert b[0][0] == 0 and b[-1][1] == n
                assert all(b[r][1] == b[r + 1][0] for r in range(world - 1))
                if world > 1:
                    unit = 4 if prec == nb.PREC_FP32 else 1
                    assert all((x[0] % (nb.SLICE_ALIGN[prec] * unit) == 0) or x[0] == n for x in b)


def test_reader_parallel_ranges_equal_serial_rules(nb, tmp_path):
    """Inputs over 8 MB are parsed in per-thread ranges cut after a line feed; a cut inside a
    record (the scanf grammar allows line feeds between the fields) or an error anywhere falls
    back to the serial parser, so the result never depends on the thread count."""
    rng = np.random.default_rng(11)
    n = 120_000
    rows = np.concatenate([rng.standard_normal((n, 6)) * 10, rng.uniform(0.5, 2, (n, 1))], axis=1)
    lines = [",".join("%.17g" % x for x in r) for r in rows]
    path = str(tmp_path / "big.csv")
    open(path, "w").write("\n".join(lines) + "\n")
    assert os.path.getsize(path) > (8 << 20)
    pos, vel = nb.read_bodies(path)
    assert np.array_equal(pos, rows[:, [0, 1, 2, 6]]) and np.array_equal(vel, rows[:, 3:6])
    # every record broken over two lines: whatever the cuts hit, the bodies are the same
    open(path, "w").write("\n".join(l.replace(",", ",\n", 3).replace(",\n", ",", 2) for l in lines))
    pos2, vel2 = nb.read_bodies(path)
    assert np.array_equal(pos2, pos) and np.array_equal(vel2, vel)
    # a malformed record in the middle is still an error
    bad = list(lines)
    bad[n // 2 + 17] = bad[n // 2 + 17].replace(",", ";", 1)
    open(path, "w").write("\n".join(bad) + "\n")
    with pytest.raises(ValueError):
        nb.read_bodies(path)
    # a truncated last record is dropped, as fscanf's short count does in the reference
    open(path, "w").write("\n".join(lines) + "\n1,2,3,")
    pos3, _ = nb.read_bodies(path)
    assert np.array_equal(pos3, pos)


def test_threaded_frame_writer_equals_serial_formatter(nb, tmp_path):
    """nbh_write_frame formats 1024-body chunks on OpenMP threads and writes them in body order:
    the file is byte for byte what the single-pass formatter produces."""
    h = nb.load_host_library()
    libc = C.CDLL(None)
    libc.fopen.restype = C.c_void_p
    libc.fopen.argtypes = [C.c_char_p, C.c_char_p]
    libc.fclose.argtypes = [C.c_void_p]
    rng = np.random.default_rng(9)
    for n in (0, 1, 1023, 1024, 1025, 40_000):
        pm = np.ascontiguousarray(np.concatenate([rng.standard_normal((n, 3)) * 100,
                                                  rng.uniform(0.1, 2, (n, 1))], axis=1))
        path = str(tmp_path / ("f%d.csv" % n))
        fp = libc.fopen(path.encode(), b"w")
        assert fp
        h.nbh_write_frame.argtypes = [C.c_void_p, C.c_size_t, np.ctypeslib.ndpointer(np.float64, flags="C")]
        assert h.nbh_write_frame(fp, n, pm) == 0
        libc.fclose(fp)
        assert open(path, "rb").read() == nb.format_frame(pm)
