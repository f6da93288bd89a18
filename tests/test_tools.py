"""tools/check_error.py: the headless restatement of utils/check_error.py:19-38 (frame-file
comparison with a regression line through body 0's |dx|).  CPU only: the frame files come
from the oracle's serial and OpenMP-ordered loops."""
import importlib.util
import json
import os
import subprocess
import sys

import numpy as np

import oracle as O
from conftest import ROOT, golden

TOOL = os.path.join(ROOT, "tools", "check_error.py")


def load_tool():
    spec = importlib.util.spec_from_file_location("check_error", TOOL)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_check_error_on_oracle_frames(oracle_lib, tmp_path):
    g = golden("20_bodies_g1")
    a, b = str(tmp_path / "a.csv"), str(tmp_path / "b.csv")
    O.propagate(oracle_lib, g["pos_xyzm"], g["vel"], 120, 0.01, frames=a)
    pm = g["pos_xyzm"].copy()
    pm[0, 0] += 1e-3                       # a perturbed run: the error of body 0 grows with time
    O.propagate(oracle_lib, pm, g["vel"], 120, 0.01, frames=b)
    ce = load_tool()
    fa, fb = ce.load_frames(a), ce.load_frames(b)
    assert fa.shape == fb.shape == (121, 20, 5)
    out, series = ce.compare(fa, fb)
    # the same numbers, computed the way the reference script does (flat rows, stride n_corps)
    d1, d2 = np.loadtxt(a, delimiter=","), np.loadtxt(b, delimiter=",")
    n_corps = int(d1[:, 0].max()) + 1
    diff = np.abs(d1 - d2)
    y = diff[[i * n_corps for i in range(len(d1) // n_corps)], :]
    assert np.array_equal(series[:, 0], y[:, 1])
    m, q = np.polyfit(np.arange(30, 100), y[30:100, 1], 1)
    assert out["fit"] == {"from": 30, "to": 100, "slope": float(m), "intercept": float(q)}
    assert out["max_abs_error"] == diff[:, 1:4].max() and out["max_abs_error"] >= 1e-3 - 1e-6
    # identical files: all zeros
    out0, _ = ce.compare(fa, fa)
    assert out0["max_abs_error"] == 0.0 and out0["fit"]["slope"] == 0.0


def test_check_error_command_line(oracle_lib, tmp_path):
    g = golden("10_bodies")
    a, b, s = (str(tmp_path / n) for n in ("a.csv", "b.csv", "s.csv"))
    O.propagate(oracle_lib, g["pos_xyzm"], g["vel"], 40, 0.01, frames=a)
    O.propagate(oracle_lib, g["pos_xyzm"], g["vel"], 40, 0.01, frames=b, threads=2)
    r = subprocess.run([sys.executable, TOOL, a, b, "--series", s, "--max-error", "1e-6"],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    out = json.loads(r.stdout)
    assert out["frames"] == 41 and out["bodies"] == 10 and out["fit"]["to"] == 41
    assert len(open(s).read().splitlines()) == 41
    pm = g["pos_xyzm"].copy()
    pm[3, 1] += 0.5
    O.propagate(oracle_lib, pm, g["vel"], 40, 0.01, frames=b)
    r = subprocess.run([sys.executable, TOOL, a, b, "--max-error", "1e-6"], capture_output=True, text=True)
    assert r.returncode == 1
