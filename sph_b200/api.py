"""ctypes binding of include/sphb.h."""
from __future__ import annotations

import ctypes as C
import os
import re
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)

ISOTHERMAL, ADIABATIC = 0, 1
ORDER_REFERENCE, ORDER_JACOBI = 0, 1
SWEEP_AUTO, SWEEP_WORKLIST, SWEEP_SERIAL = 0, 1, 2
PHASES = ("kick_drift", "sort", "ghosts", "density", "force_first", "force_refine")
FIELDS = ("mass", "pos", "vel", "acc", "h", "du_dt", "u", "density", "pressure")

# the reference's Particle, sph/basictypes.hpp:49-116 (88 bytes)
PARTICLE_DTYPE = np.dtype({
    "names": ["id", "mass", "pos", "vel", "acc", "h", "du_dt", "u", "density", "pressure", "type"],
    "formats": ["<i4"] + ["<f8"] * 9 + ["<i4"],
    "offsets": [0, 8, 16, 24, 32, 40, 48, 56, 64, 72, 80],
    "itemsize": 88,
})


class SphbConfig(C.Structure):
    # the reference's Config, sph/basictypes.hpp:21-31 (64 bytes)
    _fields_ = [("n_part", C.c_int), ("mass", C.c_double), ("pressure_calc", C.c_int),
                ("limit", C.c_double), ("v_0", C.c_double), ("h_factor", C.c_double),
                ("t_i", C.c_double), ("n_ghost", C.c_int)]


class SphbOptions(C.Structure):
    _fields_ = [("variable_h", C.c_int), ("fixed_h", C.c_double), ("order", C.c_int),
                ("max_sweeps", C.c_int), ("device", C.c_int), ("pad_capacity", C.c_int),
                ("capacity", C.c_int), ("sort_every_step", C.c_int), ("sweep_rel_tol", C.c_double),
                ("sweep_mode", C.c_int), ("sweep_group", C.c_int), ("sweep_segment", C.c_int),
                ("sweep_burn_in", C.c_int), ("reserved", C.c_int * 2)]


class SphbCounters(C.Structure):
    _fields_ = [("n_alive", C.c_int), ("n_ghost", C.c_int), ("steps", C.c_longlong),
                ("sweeps_total", C.c_longlong), ("sweeps_last", C.c_int), ("sorts", C.c_longlong),
                ("fallbacks", C.c_longlong), ("newton_hist", C.c_longlong * 12),
                ("nbr_min", C.c_int), ("nbr_max", C.c_int), ("nbr_mean", C.c_double),
                ("h_max", C.c_double), ("launches", C.c_longlong), ("force_pairs_mean", C.c_double),
                ("density_scans_mean", C.c_double), ("refine_tiles", C.c_int), ("refine_tile_evals", C.c_int),
                ("sweep_steps", C.c_longlong), ("sweep_group", C.c_int), ("sweep_segment", C.c_int),
                ("sweep_burn_in", C.c_int), ("sweep_flagged", C.c_int), ("sweep_verify_fail", C.c_int)]


assert C.sizeof(SphbConfig) == 64


@dataclass
class Config:
    """Same keys as the reference's Config / config.txt."""
    n_part: int
    mass: float
    pressure_calc: int
    limit: float = 2.0
    v_0: float = 1.0
    h_factor: float = 2.0
    t_i: float = 0.005
    n_ghost: int = 0

    def to_c(self) -> SphbConfig:
        return SphbConfig(self.n_part, self.mass, self.pressure_calc, self.limit, self.v_0,
                          self.h_factor, self.t_i, self.n_ghost)


@dataclass
class Counters:
    n_alive: int
    n_ghost: int
    steps: int
    sweeps_total: int
    sweeps_last: int
    sorts: int
    fallbacks: int
    newton_hist: list
    nbr_min: int
    nbr_max: int
    nbr_mean: float
    h_max: float
    launches: int = 0
    force_pairs_mean: float = 0.0
    density_scans_mean: float = 0.0
    refine_tiles: int = 0
    refine_tile_evals: int = 0
    sweep_steps: int = 0
    sweep_group: int = 0
    sweep_segment: int = 0
    sweep_burn_in: int = 0
    sweep_flagged: int = 0
    sweep_verify_fail: int = 0


class SphbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("sphb error %d: %s" % (code, msg))
        self.code = code
        self.msg = msg


def lib_path() -> str:
    # SPHB_LIB: an alternative build of the same library (kernel experiments, tools/)
    return os.environ.get("SPHB_LIB") or os.path.join(_HERE, "libsphb.so")


def declared_symbols() -> list:
    """Every function include/sphb.h declares."""
    text = open(os.path.join(_ROOT, "include", "sphb.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sphb_[a-z_]+)\s*\(", text)))


_lib = None


def load_library():
    """Load libsphb.so (built in-tree by sph_b200/build.py).  Fails loudly if missing."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise SphbError(12, "libsphb.so is not built (run `python -m sph_b200.build`); "
                            "there is no CPU fallback")
    L = C.CDLL(path)
    p, i, d = C.c_void_p, C.c_int, C.c_double
    ip = C.POINTER(C.c_int)
    L.sphb_default_options.argtypes = [C.POINTER(SphbOptions)]
    L.sphb_default_options.restype = None
    L.sphb_create.argtypes = [C.POINTER(SphbConfig), C.POINTER(SphbOptions), C.POINTER(p)]
    L.sphb_destroy.argtypes = [p]
    L.sphb_destroy.restype = None
    L.sphb_last_error.argtypes = [p]
    L.sphb_last_error.restype = C.c_char_p
    L.sphb_set_stream.argtypes = [p, p]
    L.sphb_upload.argtypes = [p, p, i, i]
    L.sphb_download.argtypes = [p, p, i, ip, ip, i]
    L.sphb_download_strided.argtypes = [p, p, i, i, ip, ip, i]
    L.sphb_download_fields.argtypes = [p, p, p, i, ip, ip]
    L.sphb_download_counts.argtypes = [p, p, p, i]
    L.sphb_init_lattice.argtypes = [p]
    L.sphb_init_developed.argtypes = [p, d]
    L.sphb_slab_init_developed.argtypes = [p, d]
    L.sphb_slab_init_developed_fill.argtypes = [p, d]
    L.sphb_density.argtypes = [p, i]
    L.sphb_forces.argtypes = [p]
    L.sphb_omega.argtypes = [p, p, i]
    L.sphb_ghosts.argtypes = [p, ip]
    L.sphb_kick_drift.argtypes = [p]
    L.sphb_step.argtypes = [p, i]
    L.sphb_get_counters.argtypes = [p, C.POINTER(SphbCounters)]
    L.sphb_diagnostics.argtypes = [p, C.POINTER(d)]
    L.sphb_get_config.argtypes = [p, C.POINTER(SphbConfig)]
    L.sphb_set_profiling.argtypes = [p, i]
    L.sphb_get_phase_times.argtypes = [p, C.POINTER(d), C.POINTER(C.c_longlong), i]
    ll = C.c_longlong
    L.sphb_slab_configure.argtypes = [p, i, i, i, ll, ll]
    L.sphb_slab_message_doubles.argtypes = [p, i]
    L.sphb_slab_pack.argtypes = [p, i, i, p]
    L.sphb_slab_unpack.argtypes = [p, i, i, p]
    for name in ('step_begin', 'density', 'force_first', 'refine', 'step_end', 'init_lattice', 'prepass', 'records', 'forces', 'resort', 'raise'):
        getattr(L, 'sphb_slab_' + name).argtypes = [p]
    L.sphb_slab_poll.argtypes = [p, ip]
    L.sphb_slab_agree.argtypes = [p, ip]
    L.sphb_nccl_unique_id.argtypes = [p]
    L.sphb_slab_comm_init.argtypes = [p, p]
    L.sphb_slab_init_particles.argtypes = [p]
    L.sphb_slab_step.argtypes = [p, i]
    L.sphb_slab_rounds_last.argtypes = [p]
    L.sphb_density_static.argtypes = [p, p, p, i, d, d, d, i, p, p, p]
    L.sphb_build_arch.argtypes = []
    L.sphb_build_arch.restype = C.c_char_p
    _lib = L
    return L


STATIC_MAX_PARTICLES = 1 << 20   # SPHB_STATIC_MAX_PARTICLES, include/sphb.h


def density_static(d_x, d_h, d_rho, n, mass, h_factor, limit, variable_h, d_h_out, d_rho_out, stream=0):
    """sphb_density_static: the handle-less density pass that allocates nothing.  Arguments are
    raw device addresses (ints) of n-element double arrays owned by the caller."""
    L = load_library()
    rc = L.sphb_density_static(C.c_void_p(d_x), C.c_void_p(d_h), C.c_void_p(d_rho), int(n), float(mass),
                               float(h_factor), float(limit), int(variable_h), C.c_void_p(d_h_out),
                               C.c_void_p(d_rho_out), C.c_void_p(stream))
    if rc != 0:
        raise SphbError(rc, L.sphb_last_error(None).decode())


def exported_symbols() -> list:
    L = load_library()
    return [s for s in declared_symbols() if hasattr(L, s)]


class Simulation:
    """One device-resident simulation (an sphb_handle)."""

    def __init__(self, config: Config, variable_h=True, fixed_h=0.2, order=ORDER_REFERENCE,
                 max_sweeps=64, device=-1, pad_capacity=4096, capacity=0, sort_every_step=False,
                 sweep_rel_tol=0.0, sweep_mode=SWEEP_AUTO, sweep_group=0, sweep_segment=0, sweep_burn_in=0):
        self.L = load_library()
        self.config = config
        opt = SphbOptions()
        self.L.sphb_default_options(C.byref(opt))
        opt.variable_h = int(variable_h)
        opt.fixed_h = fixed_h
        opt.order = order
        opt.max_sweeps = max_sweeps
        opt.device = device
        opt.pad_capacity = pad_capacity
        opt.capacity = capacity
        opt.sort_every_step = int(sort_every_step)
        opt.sweep_rel_tol = float(sweep_rel_tol)
        opt.sweep_mode, opt.sweep_group = int(sweep_mode), int(sweep_group)
        opt.sweep_segment, opt.sweep_burn_in = int(sweep_segment), int(sweep_burn_in)
        self.h = C.c_void_p()
        cfg = config.to_c()
        rc = self.L.sphb_create(C.byref(cfg), C.byref(opt), C.byref(self.h))
        if rc != 0:
            raise SphbError(rc, self.L.sphb_last_error(None).decode())
        self._cap = max(capacity, config.n_part - config.n_ghost) + 2 * pad_capacity

    def close(self):
        if getattr(self, "h", None):
            self.L.sphb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise SphbError(rc, self.L.sphb_last_error(self.h).decode())

    # ---- data movement
    def set_stream(self, cuda_stream_ptr):
        self._check(self.L.sphb_set_stream(self.h, C.c_void_p(cuda_stream_ptr)))

    def upload(self, particles: np.ndarray, n_ghost: int = 0):
        assert particles.dtype == PARTICLE_DTYPE and particles.flags["C_CONTIGUOUS"]
        self._check(self.L.sphb_upload(self.h, particles.ctypes.data, len(particles), n_ghost))

    def upload_fields(self, fields: np.ndarray, n_ghost: int = 0):
        """fields: [n, 9] doubles in FIELDS order."""
        n = fields.shape[0]
        rec = np.zeros(n, dtype=PARTICLE_DTYPE)
        rec["id"] = np.arange(n)
        for k, name in enumerate(FIELDS):
            rec[name] = fields[:, k]
        rec["type"][n - n_ghost:] = 1
        self.upload(rec, n_ghost)

    def download(self, out: np.ndarray = None, write_ids=True):
        if out is None:
            out = np.zeros(self._cap, dtype=PARTICLE_DTYPE)
        n, g = C.c_int(), C.c_int()
        self._check(self.L.sphb_download(self.h, out.ctypes.data, len(out), C.byref(n), C.byref(g),
                                         int(write_ids)))
        return out[: n.value], g.value

    def download_strided(self, stride, out: np.ndarray = None, write_ids=True):
        """Every stride-th alive record followed by all ghost records: (records, n_alive_written, n_ghost)."""
        if out is None:
            out = np.zeros(self._cap // stride + 2 + 2 * 4096, dtype=PARTICLE_DTYPE)
        na, g = C.c_int(), C.c_int()
        self._check(self.L.sphb_download_strided(self.h, out.ctypes.data, len(out), int(stride), C.byref(na),
                                                 C.byref(g), int(write_ids)))
        return out[: na.value + g.value], na.value, g.value

    def fields(self):
        """([n, 9] doubles, type[n], n_ghost) in array-index order."""
        f = np.zeros((self._cap, 9), dtype=np.float64)
        t = np.zeros(self._cap, dtype=np.int32)
        n, g = C.c_int(), C.c_int()
        self._check(self.L.sphb_download_fields(self.h, f.ctypes.data, t.ctypes.data, self._cap,
                                                C.byref(n), C.byref(g)))
        return f[: n.value].copy(), t[: n.value].copy(), g.value

    def counts(self):
        nw = np.zeros(self._cap, dtype=np.int32)
        nb = np.zeros(self._cap, dtype=np.int32)
        self._check(self.L.sphb_download_counts(self.h, nw.ctypes.data, nb.ctypes.data, self._cap))
        return nw, nb

    # ---- the path
    def init_lattice(self):
        self._check(self.L.sphb_init_lattice(self.h))

    def init_developed(self, t_star=0.5):
        """Synthetic developed start state (sphb_init_developed): the exact colliding-streams solution at t_star."""
        self._check(self.L.sphb_init_developed(self.h, float(t_star)))

    def density(self, include_ghosts=False):
        self._check(self.L.sphb_density(self.h, int(include_ghosts)))

    def forces(self):
        self._check(self.L.sphb_forces(self.h))

    def omega(self):
        out = np.zeros(self._cap, dtype=np.float64)
        self._check(self.L.sphb_omega(self.h, out.ctypes.data, self._cap))
        return out[: self.counters().n_alive + self.counters().n_ghost].copy()

    def ghosts(self) -> int:
        g = C.c_int()
        self._check(self.L.sphb_ghosts(self.h, C.byref(g)))
        return g.value

    def kick_drift(self):
        self._check(self.L.sphb_kick_drift(self.h))

    def step(self, nsteps=1):
        self._check(self.L.sphb_step(self.h, nsteps))

    # ---- diagnostics
    def counters(self) -> Counters:
        c = SphbCounters()
        self._check(self.L.sphb_get_counters(self.h, C.byref(c)))
        return Counters(c.n_alive, c.n_ghost, c.steps, c.sweeps_total, c.sweeps_last, c.sorts,
                        c.fallbacks, list(c.newton_hist), c.nbr_min, c.nbr_max, c.nbr_mean, c.h_max,
                        c.launches, c.force_pairs_mean, c.density_scans_mean, c.refine_tiles, c.refine_tile_evals,
                        c.sweep_steps, c.sweep_group, c.sweep_segment, c.sweep_burn_in, c.sweep_flagged,
                        c.sweep_verify_fail)

    def diagnostics(self):
        """{mass, momentum, kinetic, thermal} totals over the alive particles (sphb_diagnostics)."""
        out = (C.c_double * 4)()
        self._check(self.L.sphb_diagnostics(self.h, out))
        return dict(mass=out[0], momentum=out[1], kinetic=out[2], thermal=out[3])

    # ---- checkpoint / restart: the nine fields of every record, in array order, ARE the complete state
    def save(self, path):
        rec, ng = self.download()
        np.savez(path, records=rec, n_ghost=ng)

    def load(self, path):
        z = np.load(path)
        rec = np.ascontiguousarray(z["records"])
        self.upload(rec, int(z["n_ghost"]))

    def set_profiling(self, on=True):
        self._check(self.L.sphb_set_profiling(self.h, int(on)))

    def phase_times(self, reset=False):
        ms = (C.c_double * len(PHASES))()
        n = (C.c_longlong * len(PHASES))()
        self._check(self.L.sphb_get_phase_times(self.h, ms, n, int(reset)))
        return {p: (ms[k], n[k]) for k, p in enumerate(PHASES)}

    @property
    def n_ghost(self):
        return self.counters().n_ghost
