"""The C host programs end to end on a GPU: the reference's command line in, the
reference's frame file and stdout out.

Reference material: tests/golden/programs.npz holds md5 / line count / head / tail / stdout
of the frame file written by the reference's own exh_serial and exh_mp (-DRESULTS) for
BASELINE config 1 (`tests/1k_bodies.csv 0 1 0.01`).  The frame file has 6 decimals, the
engine agrees with the reference to ~1e-10 after 100 steps, so the files are expected to be
byte-identical; the numeric comparison (1e-6 absolute) is the contract, the md5 the bonus.
"""
import hashlib
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, golden

pytestmark = pytest.mark.gpu
BIN = os.path.join(ROOT, "n-body-simulation_b200", "bin")


def write_csv(tmp_path, name):
    """the fixture in the reference's input format (x,y,z,vx,vy,vz,mass); %.17g round-trips the
    doubles the reference's reader produced from tests/<name>.csv"""
    g = golden(name)
    p = tmp_path / (name + ".csv")
    rows = np.concatenate([g["pos_xyzm"][:, :3], g["vel"], g["pos_xyzm"][:, 3:4]], axis=1)
    p.write_text("".join(",".join("%.17g" % x for x in r) + "\n" for r in rows))
    return str(p)


def run(prog, *args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([os.path.join(BIN, prog)] + [str(a) for a in args], env=e,
                          capture_output=True, text=True, timeout=300)


def frames(path):
    return np.loadtxt(path, delimiter=",")


def test_exh_cuda_results_matches_reference_frame_file(tmp_path):
    ref = golden("programs")
    inp, out = write_csv(tmp_path, "1k_bodies"), str(tmp_path / "out.csv")
    r = run("exh_cuda_results", inp, 0, 1, 0.01, out)
    assert r.returncode == 0, r.stderr
    lines = r.stdout.splitlines()
    # reference: "Start: 0.000000, end: 1.000000, delta time: 0.010000, time steps: 100, ents: 1000"
    assert lines[0].startswith("Start: 0.000000, end: 1.000000, delta time: 0.010000, "
                               "time steps: 100, ents: 1000")
    assert lines[1] == "Completed. Output file: " + out
    assert lines[2].startswith("Elapsed wall time: ") and lines[2].endswith(" s")
    data = open(out, "rb").read()
    text = data.decode().splitlines()
    assert len(text) == int(ref["exh_serial_results_lines"]) == 101 * 1000
    # head / tail were cut by bytes: drop the broken line at the cut
    head = str(ref["exh_serial_results_head"]).splitlines()[:-1]
    tail = str(ref["exh_serial_results_tail"]).splitlines()[1:]
    assert text[:len(head)] == head
    got_tail = np.array([[float(x) for x in ln.split(",")] for ln in text[-len(tail):]])
    want_tail = np.array([[float(x) for x in ln.split(",")] for ln in tail])
    assert np.abs(got_tail - want_tail).max() <= 1e-6
    assert hashlib.md5(data).hexdigest() == str(ref["exh_serial_results_md5"])


def test_exh_cuda_five_frames_and_old_argv(tmp_path):
    """grid_size / block_size of the reference's exh_cuda are accepted and ignored"""
    ref = str(golden("programs")["frames_100_bodies_g1_5steps"]).splitlines()
    inp, out = write_csv(tmp_path, "100_bodies_g1"), str(tmp_path / "o.csv")
    r = run("exh_cuda_results", inp, 0, 0.05, 0.01, out, 4, 128)
    assert r.returncode == 0, r.stderr
    got = open(out).read().splitlines()
    assert len(got) == len(ref)
    a = np.array([[float(x) for x in ln.split(",")] for ln in got])
    b = np.array([[float(x) for x in ln.split(",")] for ln in ref])
    assert np.abs(a - b).max() <= 1e-6
    assert got == ref


def test_exh_cuda_without_results_ignores_output_name(tmp_path):
    inp, out = write_csv(tmp_path, "33_bodies"), str(tmp_path / "never.csv")
    r = run("exh_cuda", inp, 0, 0.1, 0.01, out)
    assert r.returncode == 0 and not os.path.exists(out)
    assert "time steps: 10, ents: 33" in r.stdout      # float step count (serial/exhaustive.c:53)
    r = run("exh_cuda", inp, 0, 0.1, 0.01, out, env={"NB_PRECISION": "fp32", "NB_RESULTS": "1"})
    assert r.returncode == 0 and len(open(out).read().splitlines()) == 11 * 33


def test_checkpoint_restarts_bitwise(tmp_path):
    """NB_CHECKPOINT writes the final state in INPUT format (%.17g): 6 steps == 3 + 3 steps."""
    inp = write_csv(tmp_path, "100_bodies_g1")
    a, b, c = (str(tmp_path / n) for n in ("a.csv", "b.csv", "c.csv"))
    dt = 0.015625  # 1/64: step counts are exact in the reference's float arithmetic
    assert run("exh_cuda", inp, 0, 6 * dt, dt, "x", env={"NB_CHECKPOINT": a}).returncode == 0
    assert run("exh_cuda", inp, 0, 3 * dt, dt, "x", env={"NB_CHECKPOINT": b}).returncode == 0
    assert run("exh_cuda", b, 3 * dt, 6 * dt, dt, "x", env={"NB_CHECKPOINT": c}).returncode == 0
    assert open(a).read() == open(c).read()


def test_error_exits(tmp_path):
    inp = write_csv(tmp_path, "10_bodies")
    r = run("exh_cuda", inp, 0, 1)
    assert r.returncode == 1 and "Usage" in r.stderr
    r = run("exh_cuda", str(tmp_path / "missing.csv"), 0, 1, 0.01, "o")
    assert r.returncode == 1 and "Error opening file" in r.stderr
    bad = tmp_path / "bad.csv"
    bad.write_text("1,2,3,4,5,6,7\n1,2,three,4,5,6,7\n")
    r = run("exh_cuda", str(bad), 0, 1, 0.01, "o")
    assert r.returncode == 1 and "not well formed" in r.stderr


def test_barnes_hut_cuda_program(tmp_path, oracle_lib):
    """barnes-hut-cuda: frames after every drift, no initial frame (cuda/barnes-hut.cu:1064)"""
    g = golden("100_bodies_g1")
    inp, out = write_csv(tmp_path, "100_bodies_g1"), str(tmp_path / "bh.csv")
    r = run("barnes-hut-cuda", inp, 0, 0.05, 0.01, out)
    assert r.returncode == 0, r.stderr
    f = frames(out)
    assert f.shape == (5 * 100, 5)
    pm, v = g["pos_xyzm"].copy(), g["vel"].copy()
    oracle_lib.orc_bh_propagate(100, pm, v, 5, np.float32(0.01))
    last = f[-100:]
    assert np.array_equal(last[:, 0], np.arange(100))
    assert np.abs(last[:, 1:4] - pm[:, :3]).max() <= 1e-6 and np.abs(last[:, 4] - pm[:, 3]).max() <= 1e-6


def test_barnes_hut_cuda_matches_reference_program_100_steps(tmp_path):
    """tests/golden/programs_bh.npz: the reference's serial/barnes-hut.c (-DRESULTS) on
    1k_bodies.csv `0 1 0.01`, run in the build container.  Its file starts with the initial
    frame, which the CUDA program (and ours) leaves out; the 100 frames that follow are compared
    number by number (one unit of the sixth decimal) and, since the walk reproduces the
    reference's interaction lists, byte for byte."""
    ref = golden("programs_bh")
    inp, out = write_csv(tmp_path, "1k_bodies"), str(tmp_path / "bh.csv")
    r = run("barnes-hut-cuda", inp, 0, 1, 0.01, out)
    assert r.returncode == 0, r.stderr
    assert "time steps: 100, ents: 1000" in r.stdout
    data = open(out, "rb").read()
    lines = data.decode().splitlines()
    assert len(lines) == int(ref["bh_serial_1k_lines"]) - 1000 == 100 * 1000
    want_last = str(ref["bh_serial_1k_last_frame"]).splitlines()
    got = np.array([[float(x) for x in ln.split(",")] for ln in lines[-1000:]])
    want = np.array([[float(x) for x in ln.split(",")] for ln in want_last])
    assert np.abs(got - want).max() <= 1.0000001e-6
    assert hashlib.md5(data).hexdigest() == str(ref["bh_serial_1k_steps_md5"])


def test_barnes_hut_cuda_five_frames_equal_reference(tmp_path):
    ref = str(golden("programs_bh")["bh_serial_100_bodies_g1_5steps"]).splitlines()[100:]
    inp, out = write_csv(tmp_path, "100_bodies_g1"), str(tmp_path / "bh5.csv")
    r = run("barnes-hut-cuda", inp, 0, 0.05, 0.01, out)
    assert r.returncode == 0, r.stderr
    assert open(out).read().splitlines() == ref
    # README.md:67 of the reference: an optional cache_sz_MB argument; anything else is a usage error
    r = run("barnes-hut-cuda", inp, 0, 0.05, 0.01, out, 64)
    assert r.returncode == 0 and open(out).read().splitlines() == ref
    r = run("barnes-hut-cuda", inp, 0, 0.05)
    assert r.returncode == 255 and "Usage" in r.stderr      # `return -1`, cuda/barnes-hut.cu:843


def test_periodic_checkpoints_do_not_change_the_run(tmp_path):
    """NB_CHECKPOINT_EVERY=k: the state is read back (closing the step) and written every k
    steps; the final state is bit-identical to the uninterrupted run, and a run restarted from
    an intermediate checkpoint arrives at the same state."""
    inp = write_csv(tmp_path, "100_bodies_g1")
    a, b, mid = (str(tmp_path / n) for n in ("a.csv", "b.csv", "mid.csv"))
    dt = 0.015625
    assert run("exh_cuda", inp, 0, 7 * dt, dt, "x", env={"NB_CHECKPOINT": a}).returncode == 0
    r = run("exh_cuda", inp, 0, 7 * dt, dt, "x", env={"NB_CHECKPOINT": b, "NB_CHECKPOINT_EVERY": "3"})
    assert r.returncode == 0
    assert r.stderr.count("checkpoint after step") == 2 and "after step 6 (t = 0.093750)" in r.stderr
    assert open(a).read() == open(b).read() and not os.path.exists(b + ".tmp")
    # stop at step 3 (what the first intermediate checkpoint held), restart from there
    assert run("exh_cuda", inp, 0, 3 * dt, dt, "x", env={"NB_CHECKPOINT": mid}).returncode == 0
    assert run("exh_cuda", mid, 3 * dt, 7 * dt, dt, "x", env={"NB_CHECKPOINT": mid}).returncode == 0
    assert open(mid).read() == open(a).read()


def _gpus():
    from conftest import load_package
    return load_package().device_count()


@pytest.mark.skipif(_gpus() < 2, reason="needs 2 GPUs")
def test_programs_on_two_gpus(tmp_path):
    """NB_GPUS=2: one engine per GPU inside the same program (i-slices + all-gather per step).
    Direct sum: the frame file agrees with the one-GPU run to the sixth decimal and the final
    state to 1e-12 (own-slice-first summation order); Barnes-Hut: bit for bit."""
    inp = write_csv(tmp_path, "1k_bodies")
    o1, o2, c1, c2 = (str(tmp_path / n) for n in ("o1.csv", "o2.csv", "c1.csv", "c2.csv"))
    r1 = run("exh_cuda_results", inp, 0, 0.2, 0.01, o1, env={"NB_CHECKPOINT": c1})
    r2 = run("exh_cuda_results", inp, 0, 0.2, 0.01, o2, env={"NB_CHECKPOINT": c2, "NB_GPUS": "2"})
    assert r1.returncode == 0 and r2.returncode == 0, r2.stderr
    assert "time steps: 20, ents: 1000" in r2.stdout and "per GPU" in r2.stdout
    f1, f2 = frames(o1), frames(o2)
    assert f1.shape == f2.shape == (21 * 1000, 5) and np.abs(f1 - f2).max() <= 1.0000001e-6
    s1, s2 = np.loadtxt(c1, delimiter=","), np.loadtxt(c2, delimiter=",")
    assert np.abs(s1 - s2).max() <= 1e-12 * np.abs(s1).max()
    b1, b2 = str(tmp_path / "b1.csv"), str(tmp_path / "b2.csv")
    assert run("barnes-hut-cuda", inp, 0, 0.1, 0.01, b1).returncode == 0
    r = run("barnes-hut-cuda", inp, 0, 0.1, 0.01, b2, env={"NB_GPUS": "2"})
    assert r.returncode == 0, r.stderr
    assert open(b1, "rb").read() == open(b2, "rb").read()
    # periodic checkpoints are collective: every rank reads the state back at the same steps
    r = run("exh_cuda", inp, 0, 0.07, 0.01, "x", env={"NB_GPUS": "2", "NB_CHECKPOINT": c2,
                                                         "NB_CHECKPOINT_EVERY": "3"})
    assert r.returncode == 0 and r.stderr.count("checkpoint after step") == 2


def test_energy_report_matches_get_energy(tmp_path):
    """NB_ENERGY=1: KE/PE of the initial and final state as common/get_energy.c computes them
    (golden values from the reference's get_energy on the reference's own trajectories)."""
    g = golden("1k_bodies")
    inp = write_csv(tmp_path, "1k_bodies")
    r = run("exh_cuda", inp, 0, 0.1, 0.01, "x", env={"NB_ENERGY": "1"})
    assert r.returncode == 0, r.stderr
    got = {}
    for ln in r.stderr.splitlines():
        if " KE: " in ln:
            when, rest = ln.split(" KE: ")
            ke, pe = rest.split(", PE: ")
            got[when] = (float(ke), float(pe))
    assert abs(got["initial"][0] - float(g["ke_0"])) <= 1e-12 * abs(float(g["ke_0"]))
    assert abs(got["initial"][1] - float(g["pe_0"])) <= 1e-12 * abs(float(g["pe_0"]))
    assert abs(got["final"][0] - float(g["ke_10"])) <= 1e-9 * abs(float(g["ke_10"]))
    assert abs(got["final"][1] - float(g["pe_10"])) <= 1e-9 * abs(float(g["pe_10"]))
