#!/usr/bin/env python
"""Golden output of the reference's own Barnes-Hut program (serial/barnes-hut.c built with
-DRESULTS by oracle/Makefile -> oracle/_ref/barnes_hut_results), run in THIS container on
/root/reference/tests/1k_bodies.csv `0 1 0.01` and 100_bodies_g1.csv `0 0.05 0.01`.
The serial program writes the initial frame first (serial/barnes-hut.c:460-466); the CUDA
program the engine's barnes-hut-cuda mirrors does not (cuda/barnes-hut.cu:1063-1066), so the
digests are taken over the bytes AFTER the first frame as well.

    python tests/golden/make_golden_bh_programs.py     ->  tests/golden/programs_bh.npz
"""
import hashlib
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
import oracle as O  # noqa: E402

REF = "/root/reference"


def after_first_frame(data, n):
    pos = 0
    for _ in range(n):
        pos = data.index(b"\n", pos) + 1
    return data[pos:]


def main():
    O.build()
    meta = {}
    with tempfile.TemporaryDirectory() as td:
        out = os.path.join(td, "bh1k.csv")
        r = subprocess.run([os.path.join(O.REFDIR, "barnes_hut_results"),
                            os.path.join(REF, "tests", "1k_bodies.csv"), "0", "1", "0.01", out],
                           capture_output=True, text=True, check=True)
        data = open(out, "rb").read()
        steps = after_first_frame(data, 1000)
        meta["bh_serial_1k_stdout"] = r.stdout
        meta["bh_serial_1k_lines"] = data.count(b"\n")
        meta["bh_serial_1k_md5"] = hashlib.md5(data).hexdigest()
        meta["bh_serial_1k_steps_md5"] = hashlib.md5(steps).hexdigest()
        meta["bh_serial_1k_last_frame"] = b"".join(data.splitlines(True)[-1000:]).decode()
        out = os.path.join(td, "bh100.csv")
        subprocess.run([os.path.join(O.REFDIR, "barnes_hut_results"),
                        os.path.join(REF, "tests", "100_bodies_g1.csv"), "0", "0.05", "0.01", out],
                       capture_output=True, check=True)
        meta["bh_serial_100_bodies_g1_5steps"] = open(out, "rb").read().decode()
    np.savez_compressed(os.path.join(HERE, "programs_bh.npz"),
                        **{k: np.str_(v) if isinstance(v, str) else v for k, v in meta.items()})
    print(meta["bh_serial_1k_md5"], meta["bh_serial_1k_steps_md5"], meta["bh_serial_1k_lines"])
    print(meta["bh_serial_1k_stdout"])


if __name__ == "__main__":
    main()
