#!/usr/bin/env python
"""Generate tests/golden/*.npz from the REFERENCE ITSELF (run in the build container).

Inputs  : the reference's body fixtures, read by the reference's own get_entities()
          (/root/reference/tests/*.csv through oracle/_ref/libref_exh_serial.so).
Outputs : what the reference's own compiled code computes from them, at full
          double precision (the reference's output file keeps 6 decimals only):
            acc0            acceleration() on the initial state  (serial/exhaustive.c:155)
            pos_K / vel_K   state after K steps of propagation() (serial/exhaustive.c:192), dt = 0.01f
            ke_K / pe_K     get_energy() (common/get_energy.c:7) at step 0 and K
            bh_acc0         serial Barnes-Hut accelerations (serial/barnes-hut.c:470-473)
            frames_md5      md5 of the -DRESULTS frame file of exh_serial for `0 1 0.01`
/root/reference does not exist on the GPU box, so the tests read only the npz files.

    python tests/golden/make_golden.py
"""
import ctypes as C
import hashlib
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import oracle as O  # noqa: E402

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))
DT = np.float32(0.01)

CASES = {  # name -> step counts to store
    "10_bodies": (1, 10, 100),
    "20_bodies_g1": (1, 10, 100),
    "33_bodies": (1, 10, 100),
    "100_bodies_g1": (1, 10, 100),
    "500_bodies_g1": (1, 10),
    "1k_bodies": (1, 10, 100),
    "sun_earth": (1, 10),
}


def energy(lib, pm, v, G=1.0):
    ke, pe = C.c_double(), C.c_double()
    pos = np.ascontiguousarray(pm[:, :3])
    mass = np.ascontiguousarray(pm[:, 3])
    lib.ref_get_energy(pos, np.ascontiguousarray(v), mass, G, pm.shape[0], C.byref(ke), C.byref(pe))
    return ke.value, pe.value


def main():
    O.build(ref=True)
    rs = O.load_ref("exh_serial")
    re_ = O.load_ref("energy")
    rb = O.load_ref("bh_serial")

    for name, ks in CASES.items():
        path = os.path.join(REF, "tests", name + ".csv")
        _, pm, v = O.read_bodies(rs, path, "ref_read_bodies")
        n = pm.shape[0]
        d = {"pos_xyzm": pm, "vel": v, "dt": DT, "csv_text": np.bytes_(open(path, "rb").read())
             if name == "sun_earth" else np.bytes_(b"")}
        d["acc0"] = O.acceleration(rs, pm, "ref_acceleration")
        d["ke_0"], d["pe_0"] = energy(re_, pm, v)
        for k in ks:
            p2, v2 = pm.copy(), v.copy()
            rs.ref_propagate(n, p2, v2, k, DT, b"/dev/null")
            d["pos_%d" % k], d["vel_%d" % k] = p2, v2
            d["ke_%d" % k], d["pe_%d" % k] = energy(re_, p2, v2)
        if n >= 2 and name != "sun_earth":
            bh = np.zeros((n, 3))
            st = np.zeros(3)
            rb.ref_bh_acceleration(n, pm, bh, st.ctypes.data)
            d["bh_acc0"], d["bh_stats"] = bh, st
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
        print(name, n, "ok")

    # 10k: input, one acceleration, K = 10, BH accelerations
    path = os.path.join(REF, "tests", "10k_bodies_g1.csv")
    _, pm, v = O.read_bodies(rs, path, "ref_read_bodies")
    ro = O.load_ref("exh_omp")
    ro.ref_set_threads(os.cpu_count())
    d = {"pos_xyzm": pm, "vel": v, "dt": DT}
    d["acc0"] = O.acceleration(ro, pm, "ref_acceleration")
    p2, v2 = pm.copy(), v.copy()
    ro.ref_propagate(pm.shape[0], p2, v2, 10, DT, b"/dev/null")
    d["pos_10"], d["vel_10"] = p2, v2
    d["ke_0"], d["pe_0"] = energy(re_, pm, v)
    d["ke_10"], d["pe_10"] = energy(re_, p2, v2)
    bh = np.zeros_like(d["acc0"])
    st = np.zeros(3)
    rb.ref_bh_acceleration(pm.shape[0], pm, bh, st.ctypes.data)
    d["bh_acc0"], d["bh_stats"] = bh, st
    np.savez_compressed(os.path.join(OUT, "10k_bodies_g1.npz"), **d)
    print("10k ok")

    # the reference's stand-alone programs: frame file + stdout for the headline CPU config
    meta = {}
    with tempfile.TemporaryDirectory() as td:
        for prog in ("exh_serial_results", "exh_mp_results"):
            out = os.path.join(td, prog + ".csv")
            cmd = [os.path.join(O.REFDIR, prog), os.path.join(REF, "tests", "1k_bodies.csv"),
                   "0", "1", "0.01", out] + (["4"] if "mp" in prog else [])
            r = subprocess.run(cmd, capture_output=True, text=True, check=True)
            data = open(out, "rb").read()
            meta[prog + "_md5"] = hashlib.md5(data).hexdigest()
            meta[prog + "_lines"] = data.count(b"\n")
            meta[prog + "_head"] = data[:400].decode()
            meta[prog + "_tail"] = data[-400:].decode()
            meta[prog + "_stdout"] = r.stdout
        # small case kept in full for byte-for-byte frame-writer tests
        out = os.path.join(td, "f100.csv")
        subprocess.run([os.path.join(O.REFDIR, "exh_serial_results"),
                        os.path.join(REF, "tests", "100_bodies_g1.csv"), "0", "0.05", "0.01", out],
                       capture_output=True, check=True)
        meta["frames_100_bodies_g1_5steps"] = open(out, "rb").read().decode()
        r = subprocess.run([os.path.join(O.REFDIR, "get_energy_kat")], capture_output=True, text=True)
        meta["get_energy_kat_stdout"] = r.stdout
    np.savez_compressed(os.path.join(OUT, "programs.npz"), **{k: np.str_(v) if isinstance(v, str) else v
                                                              for k, v in meta.items()})
    print(meta["exh_serial_results_md5"], meta["exh_mp_results_md5"], meta["get_energy_kat_stdout"])


if __name__ == "__main__":
    main()
