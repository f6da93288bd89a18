"""ctypes loaders for the oracle -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module.  The product package never does.

* ``load_oracle()``  -> the CPU restatement (oracle/nbody_oracle.c, bh_oracle.c)
* ``load_ref(name)`` -> the reference's own sources compiled in place into
  oracle/_ref/ (present in the build container; travels prebuilt to the GPU box)
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
BUILD = os.path.join(HERE, "_build")
REFDIR = os.path.join(HERE, "_ref")

_dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")


def build(ref=True):
    """Compile the restatement (always) and oracle/_ref (when /root/reference exists)."""
    targets = ["oracle"] + (["ref"] if ref else [])
    subprocess.run(["make", "-s", "-C", HERE] + targets, check=True,
                   stdout=subprocess.DEVNULL)


def have_ref():
    return os.path.exists(os.path.join(REFDIR, "libref_exh_serial.so"))


def load_oracle(fast=False):
    path = os.path.join(BUILD, "liboracle_fast.so" if fast else "liboracle.so")
    if not os.path.exists(path):
        build(ref=False)
    lib = C.CDLL(path)
    lib.orc_read_bodies.argtypes = [C.c_char_p, C.POINTER(C.c_uint),
                                    C.POINTER(C.POINTER(C.c_double)),
                                    C.POINTER(C.POINTER(C.c_double))]
    lib.orc_read_bodies.restype = C.c_int
    lib.orc_free.argtypes = [C.c_void_p]
    lib.orc_parse_times.argtypes = [C.c_char_p] * 3 + [C.POINTER(C.c_float)] * 3
    lib.orc_parse_times.restype = C.c_int
    lib.orc_acceleration.argtypes = [C.c_uint, _dp, _dp]
    lib.orc_acceleration_mt.argtypes = [C.c_uint, _dp, _dp, C.c_int]
    lib.orc_acceleration_f32.argtypes = [C.c_uint, _dp, _dp]
    lib.orc_acceleration_slice.argtypes = [C.c_uint, _dp, C.c_uint, C.c_uint, _dp]
    lib.orc_propagate.argtypes = [C.c_uint, _dp, _dp, C.c_int, C.c_float, C.c_char_p, C.c_int]
    lib.orc_propagate.restype = C.c_int
    lib.orc_get_energy.argtypes = [C.c_uint, _dp, _dp, C.c_double,
                                   C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.orc_get_energy_flat.argtypes = [_dp, _dp, _dp, C.c_double, C.c_int,
                                        C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.orc_bh_bounding_box.argtypes = [C.c_uint, _dp]
    lib.orc_bh_bounding_box.restype = C.c_double
    lib.orc_bh_acceleration.argtypes = [C.c_uint, _dp, _dp, C.c_void_p]
    lib.orc_bh_propagate.argtypes = [C.c_uint, _dp, _dp, C.c_int, C.c_float]
    lib.orc_bh_propagate.restype = C.c_int
    return lib


def load_ref(name):
    """name in {exh_serial, exh_omp, exh_omp_fast, energy, bh_serial}."""
    lib = C.CDLL(os.path.join(REFDIR, "libref_%s.so" % name))
    if name.startswith("exh"):
        lib.ref_set_threads.argtypes = [C.c_int]
        lib.ref_read_bodies.argtypes = [C.c_char_p, C.POINTER(C.c_uint),
                                        C.POINTER(C.POINTER(C.c_double)),
                                        C.POINTER(C.POINTER(C.c_double))]
        lib.ref_free.argtypes = [C.c_void_p]
        lib.ref_acceleration.argtypes = [C.c_uint, _dp, _dp]
        lib.ref_propagate.argtypes = [C.c_uint, _dp, _dp, C.c_int, C.c_float, C.c_char_p]
    elif name == "energy":
        lib.ref_get_energy.argtypes = [_dp, _dp, _dp, C.c_double, C.c_int,
                                       C.POINTER(C.c_double), C.POINTER(C.c_double)]
    elif name == "bh_serial":
        lib.ref_bh_acceleration.argtypes = [C.c_int, _dp, _dp, C.c_void_p]
        lib.ref_bh_nodes.argtypes = [C.c_int, _dp, _dp, C.c_int]
        lib.ref_bh_nodes.restype = C.c_int
        lib.ref_bh_propagate.argtypes = [C.c_int, _dp, _dp, C.c_int, C.c_float, C.c_char_p]
    return lib


# ---------------------------------------------------------------- helpers ---
def read_bodies(lib, path, fn="orc_read_bodies"):
    """Returns (rc, pos_xyzm[n,4], vel[n,3]) through the C reader under test."""
    n = C.c_uint(0)
    p = C.POINTER(C.c_double)()
    v = C.POINTER(C.c_double)()
    rc = getattr(lib, fn)(path.encode(), C.byref(n), C.byref(p), C.byref(v))
    if rc != 0:
        return rc, None, None
    pos = np.ctypeslib.as_array(p, shape=(n.value, 4)).copy() if n.value else np.zeros((0, 4))
    vel = np.ctypeslib.as_array(v, shape=(n.value, 3)).copy() if n.value else np.zeros((0, 3))
    free = lib.orc_free if fn.startswith("orc") else lib.ref_free
    free(p)
    free(v)
    return 0, pos, vel


def acceleration(lib, pos_xyzm, fn="orc_acceleration", *extra):
    pm = np.ascontiguousarray(pos_xyzm, dtype=np.float64)
    acc = np.zeros((pm.shape[0], 3))
    getattr(lib, fn)(pm.shape[0], pm, acc, *extra)
    return acc


def propagate(lib, pos_xyzm, vel, n_steps, dt, frames=None, threads=1):
    pm = np.array(pos_xyzm, dtype=np.float64, order="C")
    v = np.array(vel, dtype=np.float64, order="C")
    rc = lib.orc_propagate(pm.shape[0], pm, v, n_steps, np.float32(dt),
                           frames.encode() if frames else None, threads)
    assert rc == 0
    return pm, v


def get_energy(lib, pos_xyzm, vel, G=1.0):
    ke, pe = C.c_double(), C.c_double()
    pm = np.ascontiguousarray(pos_xyzm, dtype=np.float64)
    v = np.ascontiguousarray(vel, dtype=np.float64)
    lib.orc_get_energy(pm.shape[0], pm, v, G, C.byref(ke), C.byref(pe))
    return ke.value, pe.value
