"""Python host over the C ABI of libnbody_b200.so (include/nbody_engine.h).

This is a thin ctypes binding plus a mirror of the reference programs'
interface (``read_bodies`` ~ get_entities, ``propagation`` ~ propagation(),
``acceleration`` ~ acceleration(); serial/exhaustive.c:110,155,192) for the
parity tests and the benchmark.  It computes nothing itself: every number
comes from the CUDA library.  If the library is missing the import of
``Engine`` fails loudly -- there is no CPU path here.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NBODY_B200_LIB") or os.path.join(HERE, "libnbody_b200.so")  # override: A/B builds
HOST_LIB_PATH = os.path.join(HERE, "libnbhost.so")

PREC_FP64, PREC_FP32 = 0, 1
ALGO_DIRECT, ALGO_BARNES_HUT = 0, 1
COMM_ID_BYTES = 128
TUNE_SCALAR_FP32, TUNE_NO_SPLIT, TUNE_NO_OVERLAP, TUNE_COUNT_VISITS, TUNE_NO_UNIFORM = 1, 2, 4, 8, 16
TUNE_NO_REORDER = 32
TUNE_NO_INTERLEAVE = 64
TUNE_NO_GRAPH = 128
TUNE_REPLICATED_SORT = 1 << 28
TUNE_EMULATE_RUNS_SHIFT = 24  # Barnes-Hut on one rank: sort as (flags >> 24 & 15) slices + merge
SLICE_ALIGN = {PREC_FP32: 64, PREC_FP64: 128}  # units; nb_direct.cuh kSliceAlign*

# every symbol include/nbody_engine.h declares
ABI_SYMBOLS = ["nb_config_default", "nb_init", "nb_upload", "nb_step", "nb_gather",
               "nb_io_slice", "nb_upload_slice", "nb_gather_slice",
               "nb_frame_begin", "nb_frame_wait", "nb_host_alloc", "nb_host_free",
               "nb_accelerations", "nb_energy", "nb_energy_static", "nb_sync", "nb_get_stats", "nb_tree_stats", "nb_free",
               "nb_comm_id", "nb_strerror", "nb_last_error", "nb_device_count", "nb_version"]
HOST_SYMBOLS = ["nbh_read_bodies", "nbh_parse_bodies", "nbh_free", "nbh_parse_times",
                "nbh_write_frame", "nbh_format_frame", "nbh_format_lf", "nbh_write_bodies"]


class NbConfig(C.Structure):
    _fields_ = [("precision", C.c_int), ("algorithm", C.c_int), ("dt", C.c_double),
                ("eps2", C.c_double), ("G", C.c_double), ("theta", C.c_double),
                ("device", C.c_int), ("rank", C.c_int), ("world", C.c_int),
                ("comm_id", C.c_void_p), ("tune_ctas_per_sm", C.c_int), ("tune_flags", C.c_int)]


class NbStats(C.Structure):
    _fields_ = [("last_step_ms", C.c_double), ("last_force_ms", C.c_double),
                ("force_launches", C.c_ulonglong), ("kernel_launches", C.c_ulonglong),
                ("interactions", C.c_ulonglong), ("steps_done", C.c_ulonglong),
                ("grid", C.c_int), ("block", C.c_int), ("smem_bytes", C.c_int),
                ("n", C.c_size_t), ("n_padded", C.c_size_t), ("slice_begin", C.c_size_t),
                ("slice_end", C.c_size_t), ("collectives", C.c_ulonglong),
                ("split_evaluations", C.c_ulonglong), ("uniform_mass", C.c_int),
                ("kernel_variant", C.c_int)]


class NbTreeInfo(C.Structure):
    _fields_ = [("nodes", C.c_uint), ("depth", C.c_uint), ("root_border", C.c_double),
                ("interactions", C.c_ulonglong), ("node_visits", C.c_ulonglong),
                ("pool_overflow", C.c_int), ("stack_overflow", C.c_int),
                ("child_tests", C.c_ulonglong), ("stage_ms", C.c_double * 8)]


class NbodyError(RuntimeError):
    def __init__(self, code, what, detail):
        super().__init__("%s failed: %s [%d] %s" % (what, detail[0], code, detail[1]))
        self.code = code


_dp = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_lib = None
_hostlib = None


def load_library():
    """dlopen libnbody_b200.so; raises if it was not built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("%s is missing: run `make` (or __graft_entry__.build()); "
                          "this package has no CPU fallback" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    lib.nb_config_default.argtypes = [C.POINTER(NbConfig)]
    lib.nb_config_default.restype = None
    lib.nb_init.argtypes = [C.POINTER(C.c_void_p), C.POINTER(NbConfig), C.c_size_t,
                            C.c_void_p, C.c_void_p]
    lib.nb_upload.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.nb_io_slice.argtypes = [C.c_void_p, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]
    lib.nb_upload_slice.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.nb_gather_slice.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.nb_step.argtypes = [C.c_void_p, C.c_int]
    lib.nb_gather.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.nb_frame_begin.argtypes = [C.c_void_p, C.c_void_p]
    lib.nb_frame_wait.argtypes = [C.c_void_p, C.c_void_p]
    lib.nb_host_alloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
    lib.nb_host_free.argtypes = [C.c_void_p]
    lib.nb_accelerations.argtypes = [C.c_void_p, C.c_void_p]
    lib.nb_energy.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.nb_energy_static.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_double,
                                     C.c_void_p, C.c_void_p]
    lib.nb_sync.argtypes = [C.c_void_p]
    lib.nb_get_stats.argtypes = [C.c_void_p, C.POINTER(NbStats)]
    lib.nb_tree_stats.argtypes = [C.c_void_p, C.POINTER(NbTreeInfo)]
    lib.nb_free.argtypes = [C.c_void_p]
    lib.nb_free.restype = None
    lib.nb_comm_id.argtypes = [C.c_void_p, C.c_size_t]
    lib.nb_strerror.argtypes = [C.c_int]
    lib.nb_strerror.restype = C.c_char_p
    lib.nb_last_error.restype = C.c_char_p
    lib.nb_device_count.restype = C.c_int
    lib.nb_version.restype = C.c_char_p
    _lib = lib
    return lib


def load_host_library():
    """The C host I/O routines (CSV reader, frame writer) -- no CUDA inside."""
    global _hostlib
    if _hostlib is not None:
        return _hostlib
    if not os.path.exists(HOST_LIB_PATH):
        raise ImportError("%s is missing: run `make`" % HOST_LIB_PATH)
    h = C.CDLL(HOST_LIB_PATH)
    h.nbh_read_bodies.argtypes = [C.c_char_p, C.POINTER(C.c_size_t),
                                  C.POINTER(C.POINTER(C.c_double)), C.POINTER(C.POINTER(C.c_double))]
    h.nbh_parse_bodies.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t),
                                   C.POINTER(C.POINTER(C.c_double)), C.POINTER(C.POINTER(C.c_double))]
    h.nbh_free.argtypes = [C.c_void_p]
    h.nbh_parse_times.argtypes = [C.c_char_p] * 3 + [C.POINTER(C.c_float)] * 3
    h.nbh_format_frame.argtypes = [C.c_char_p, C.c_size_t, C.c_size_t, _dp]
    h.nbh_format_frame.restype = C.c_size_t
    h.nbh_format_lf.argtypes = [C.c_char_p, C.c_double]
    h.nbh_format_lf.restype = C.c_size_t
    h.nbh_write_bodies.argtypes = [C.c_char_p, C.c_size_t, _dp, _dp]
    _hostlib = h
    return h


def _check(lib, rc, what):
    if rc != 0:
        raise NbodyError(rc, what, (lib.nb_strerror(rc).decode(), lib.nb_last_error().decode()))


ENERGY_STATIC_MAX_BODIES = 1048576


def energy_static(d_pos_ptr, d_vel_ptr, n, d_ke_pe_ptr, G=1.0, device=0, stream=None):
    """nb_energy_static: DEVICE pointers (ints) in the host packing, result written to the device
    pointer d_ke_pe_ptr on `stream`; allocates nothing and does not synchronise."""
    lib = load_library()
    _check(lib, lib.nb_energy_static(int(device), d_pos_ptr, d_vel_ptr, int(n), float(G),
                                     d_ke_pe_ptr, stream), "nb_energy_static")


def device_count():
    return load_library().nb_device_count()


def comm_id():
    """NB_COMM_ID_BYTES bytes for cfg.comm_id (rank 0 creates, the host distributes)."""
    lib = load_library()
    buf = C.create_string_buffer(COMM_ID_BYTES)
    _check(lib, lib.nb_comm_id(buf, COMM_ID_BYTES), "nb_comm_id")
    return buf.raw


def share_comm_id(dist, rank, device="cpu", make=None):
    """Rank 0 creates the engine communicator id, torch.distributed carries it to the others
    (the host-side half of the multi-GPU plumbing; works over gloo and nccl)."""
    import torch
    idt = torch.zeros(COMM_ID_BYTES, dtype=torch.uint8, device=device)
    if rank == 0:
        raw = (make or comm_id)()
        idt.copy_(torch.frombuffer(bytearray(raw), dtype=torch.uint8))
    dist.broadcast(idt, 0)
    return bytes(idt.cpu().numpy().tobytes())


def slice_bounds(n, world, rank, precision):
    """The engine's i-slice ownership rule, mirrored for the host side (bodies).

    Units are quads (FP32) or bodies (FP64); every rank owns the same number of
    units, a multiple of the j-tile size when world > 1 (64 quads / 128 bodies),
    the tail padded with massless bodies.  Returns (begin, end) clipped to n."""
    unit = 4 if precision == PREC_FP32 else 1
    units = (n + unit - 1) // unit
    per = (units + world - 1) // world
    if world > 1:
        a = SLICE_ALIGN[precision]
        per = (per + a - 1) // a * a
    per = max(per, 1)
    b, e = per * rank * unit, per * (rank + 1) * unit
    return min(b, n), min(e, n)


class Engine:
    """One rank of the engine: nb_init ... nb_free."""

    def __init__(self, pos_xyzm, vel_xyz, precision=PREC_FP64, dt=np.float32(0.01), eps2=0.01,
                 G=1.0, theta=0.5, algorithm=ALGO_DIRECT, device=0, rank=0, world=1,
                 comm_id_bytes=None, ctas_per_sm=0, tune_flags=0):
        self.lib = load_library()
        pos = np.ascontiguousarray(pos_xyzm, dtype=np.float64).reshape(-1, 4)
        vel = np.ascontiguousarray(vel_xyz, dtype=np.float64).reshape(-1, 3)
        if pos.shape[0] != vel.shape[0]:
            raise ValueError("positions and velocities disagree on the body count")
        self.n = pos.shape[0]
        cfg = NbConfig()
        self.lib.nb_config_default(C.byref(cfg))
        cfg.precision, cfg.algorithm = precision, algorithm
        cfg.dt = float(np.float32(dt))  # the reference's dt is a float (serial/exhaustive.c:43)
        cfg.eps2, cfg.G, cfg.theta = eps2, G, theta
        cfg.device, cfg.rank, cfg.world = device, rank, world
        cfg.tune_ctas_per_sm, cfg.tune_flags = ctas_per_sm, tune_flags
        self._idbuf = None
        if world > 1:
            self._idbuf = C.create_string_buffer(comm_id_bytes, COMM_ID_BYTES)
            cfg.comm_id = C.cast(self._idbuf, C.c_void_p)
        self.cfg = cfg
        self.handle = C.c_void_p()
        self._pinned = []
        _check(self.lib, self.lib.nb_init(C.byref(self.handle), C.byref(cfg), self.n,
                                          pos.ctypes.data, vel.ctypes.data), "nb_init")

    def upload(self, pos_xyzm, vel_xyz):
        pos = np.ascontiguousarray(pos_xyzm, dtype=np.float64)
        vel = np.ascontiguousarray(vel_xyz, dtype=np.float64)
        _check(self.lib, self.lib.nb_upload(self.handle, pos.ctypes.data, vel.ctypes.data), "nb_upload")

    def upload_raw(self, pos_ptr, vel_ptr):
        _check(self.lib, self.lib.nb_upload(self.handle, pos_ptr, vel_ptr), "nb_upload")

    def io_slice(self):
        """(begin, count): the rows of the caller's order this rank moves in upload_slice / gather_slice"""
        b, c = C.c_size_t(), C.c_size_t()
        _check(self.lib, self.lib.nb_io_slice(self.handle, C.byref(b), C.byref(c)), "nb_io_slice")
        return b.value, c.value

    def upload_slice(self, pos_rows, vel_rows):
        pos = np.ascontiguousarray(pos_rows, dtype=np.float64)
        vel = np.ascontiguousarray(vel_rows, dtype=np.float64)
        _check(self.lib, self.lib.nb_upload_slice(self.handle, pos.ctypes.data, vel.ctypes.data),
               "nb_upload_slice")

    def upload_slice_raw(self, pos_ptr, vel_ptr):
        _check(self.lib, self.lib.nb_upload_slice(self.handle, pos_ptr, vel_ptr), "nb_upload_slice")

    def gather_slice(self, with_velocities=True):
        _b, c = self.io_slice()
        pos = np.empty((c, 4))
        vel = np.empty((c, 3)) if with_velocities else None
        _check(self.lib, self.lib.nb_gather_slice(self.handle, pos.ctypes.data,
                                                  vel.ctypes.data if with_velocities else None),
               "nb_gather_slice")
        return pos, vel

    def gather_slice_raw(self, pos_ptr, vel_ptr):
        _check(self.lib, self.lib.nb_gather_slice(self.handle, pos_ptr, vel_ptr), "nb_gather_slice")

    def step(self, n_steps=1):
        _check(self.lib, self.lib.nb_step(self.handle, int(n_steps)), "nb_step")

    def gather(self, with_velocities=True):
        pos = np.empty((self.n, 4))
        vel = np.empty((self.n, 3)) if with_velocities else None
        _check(self.lib, self.lib.nb_gather(self.handle, pos.ctypes.data,
                                            vel.ctypes.data if with_velocities else None), "nb_gather")
        return pos, vel

    def frame_buffer(self):
        """a page-locked (n, 4) float64 array for frame_begin (nb_host_alloc); freed with the engine"""
        p = C.c_void_p()
        _check(self.lib, self.lib.nb_host_alloc(C.byref(p), max(self.n, 1) * 32), "nb_host_alloc")
        self._pinned.append(p)
        buf = (C.c_double * (self.n * 4)).from_address(p.value)
        return np.frombuffer(buf, dtype=np.float64).reshape(self.n, 4)

    def frame_begin(self, out):
        """starts the asynchronous read-back of the current positions into `out` ((n, 4) float64)"""
        assert out.dtype == np.float64 and out.flags.c_contiguous and out.size == self.n * 4
        _check(self.lib, self.lib.nb_frame_begin(self.handle, out.ctypes.data), "nb_frame_begin")

    def frame_wait(self, out=None):
        _check(self.lib, self.lib.nb_frame_wait(self.handle, out.ctypes.data if out is not None else None),
               "nb_frame_wait")

    def gather_raw(self, pos_ptr, vel_ptr):
        _check(self.lib, self.lib.nb_gather(self.handle, pos_ptr, vel_ptr), "nb_gather")

    def accelerations(self):
        acc = np.empty((self.n, 3))
        _check(self.lib, self.lib.nb_accelerations(self.handle, acc.ctypes.data), "nb_accelerations")
        return acc

    def energy(self):
        ke, pe = C.c_double(), C.c_double()
        _check(self.lib, self.lib.nb_energy(self.handle, C.byref(ke), C.byref(pe)), "nb_energy")
        return ke.value, pe.value

    def sync(self):
        _check(self.lib, self.lib.nb_sync(self.handle), "nb_sync")

    def stats(self):
        st = NbStats()
        _check(self.lib, self.lib.nb_get_stats(self.handle, C.byref(st)), "nb_get_stats")
        return st

    def tree_stats(self):
        ti = NbTreeInfo()
        _check(self.lib, self.lib.nb_tree_stats(self.handle, C.byref(ti)), "nb_tree_stats")
        return ti

    def close(self):
        if self.handle:
            self.lib.nb_free(self.handle)
            self.handle = C.c_void_p()
            for p in self._pinned:  # arrays from frame_buffer() die with the engine
                self.lib.nb_host_free(p)
            self._pinned = []

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---------------------------------------------------------------------------
# Mirror of the reference programs' interface (same names / argument meaning).
def read_bodies(path):
    """get_entities (cuda/exhaustive.cu:228-263): returns (pos_xyzm[n,4], vel[n,3]).
    Raises OSError(1) / ValueError like the reference's exit(1) paths."""
    h = load_host_library()
    n = C.c_size_t(0)
    p, v = C.POINTER(C.c_double)(), C.POINTER(C.c_double)()
    rc = h.nbh_read_bodies(os.fsencode(path), C.byref(n), C.byref(p), C.byref(v))
    if rc == 1:
        raise OSError(1, "Error opening file '%s'" % path)
    if rc == 3:
        raise ValueError("Error while reading file '%s': file is not well formed" % path)
    if rc == 2:
        raise MemoryError("Error during memory allocation")
    pos = np.ctypeslib.as_array(p, shape=(n.value, 4)).copy() if n.value else np.zeros((0, 4))
    vel = np.ctypeslib.as_array(v, shape=(n.value, 3)).copy() if n.value else np.zeros((0, 3))
    h.nbh_free(p)
    h.nbh_free(v)
    return pos, vel


def parse_times(start, end, dt):
    """strtof x3 and the float step count (serial/exhaustive.c:49-53)."""
    h = load_host_library()
    s, e, d = C.c_float(), C.c_float(), C.c_float()
    n = h.nbh_parse_times(str(start).encode(), str(end).encode(), str(dt).encode(),
                          C.byref(s), C.byref(e), C.byref(d))
    return s.value, e.value, d.value, n


def format_frame(pos_xyzm):
    h = load_host_library()
    pm = np.ascontiguousarray(pos_xyzm, dtype=np.float64)
    cap = pm.shape[0] * 160 + 4096
    buf = C.create_string_buffer(cap)
    k = h.nbh_format_frame(buf, cap, pm.shape[0], pm)
    return buf.raw[:k]


def acceleration(pos_xyzm, precision=PREC_FP64, **kw):
    """acceleration() of the reference (serial/exhaustive.c:155-181) on the GPU."""
    pm = np.ascontiguousarray(pos_xyzm, dtype=np.float64)
    with Engine(pm, np.zeros((pm.shape[0], 3)), precision=precision, **kw) as e:
        return e.accelerations()


def propagation(pos_xyzm, vel_xyz, n_steps, dt, output=None, precision=PREC_FP64, **kw):
    """propagation() of the reference (serial/exhaustive.c:192-255): advances
    n_steps, optionally writing the RESULTS frame file; returns (pos, vel)."""
    with Engine(pos_xyzm, vel_xyz, precision=precision, dt=dt, **kw) as e:
        if output is None:
            e.step(n_steps)
        else:
            with open(output, "wb") as fp:
                fp.write(format_frame(pos_xyzm))
                for _ in range(n_steps):
                    e.step(1)
                    p, _v = e.gather(with_velocities=False)
                    fp.write(format_frame(p))
        return e.gather()
