"""oracle/sph_oracle.py — TEST INFRASTRUCTURE, not product code.

ctypes view of oracle/libsph_oracle.so (the plain-C restatement in
oracle/sph_oracle.c).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this; the product
(sph_b200/) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libsph_oracle.so")
FIELDS = ("mass", "pos", "vel", "acc", "h", "du_dt", "u", "density", "pressure")
ISOTHERMAL, ADIABATIC = 0, 1
ORDER_REFERENCE, ORDER_JACOBI = 0, 1


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "sph_oracle.c")
    hdr = os.path.join(_HERE, "sph_oracle.h")
    stale = (not os.path.exists(_LIB)
             or os.path.getmtime(_LIB) < max(os.path.getmtime(src), os.path.getmtime(hdr)))
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "oracle"], stdout=subprocess.DEVNULL)
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    build()
    L = C.CDLL(_LIB)
    d, i, p = C.c_double, C.c_int, C.c_void_p
    L.orc_create.restype = p
    L.orc_create.argtypes = [i, d, i, d, d, d, d, i, d]
    L.orc_destroy.argtypes = [p]
    L.orc_set_order.argtypes = [p, i]
    L.orc_set_index_order_sums.argtypes = [p, i]
    L.orc_init_particles.argtypes = [p]
    L.orc_set_state.argtypes = [p, i, i, p]
    L.orc_get_state.argtypes = [p, p, p]
    L.orc_n_part.argtypes = [p]
    L.orc_n_ghost.argtypes = [p]
    L.orc_ghosts.argtypes = [p]
    L.orc_density.argtypes = [p, i, i]
    L.orc_forces.argtypes = [p, i, i]
    L.orc_kick_drift.argtypes = [p]
    L.orc_omega.argtypes = [p, p]
    L.orc_step.argtypes = [p, i]
    L.orc_get_newton_iters.argtypes = [p, p]
    L.orc_get_neighbour_counts.argtypes = [p, p]
    L.orc_fallback_count.argtypes = [p]
    L.orc_fallback_count.restype = C.c_long
    L.orc_mean_force_neighbours.argtypes = [p]
    L.orc_mean_force_neighbours.restype = d
    L.orc_last_error.argtypes = [p]
    L.orc_last_error.restype = C.c_char_p
    L.orc_w.argtypes = [d]
    L.orc_w.restype = d
    L.orc_dw.argtypes = [d]
    L.orc_dw.restype = d
    _lib = L
    return L


class OracleError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("oracle error %d: %s" % (code, msg))
        self.code = code


class Oracle:
    """CPU oracle simulation; mirrors the reference's Config keys."""

    def __init__(self, n_part, mass, pressure_calc, limit=2.0, v_0=1.0, h_factor=2.0, t_i=0.005,
                 variable_h=True, constant_h=0.2, order=ORDER_REFERENCE, index_order_sums=False,
                 init=True):
        self.L = lib()
        self.s = self.L.orc_create(n_part, mass, pressure_calc, limit, v_0, h_factor, t_i,
                                   int(variable_h), constant_h)
        self.L.orc_set_order(self.s, order)
        self.L.orc_set_index_order_sums(self.s, int(index_order_sums))
        if init:
            self._check(self.L.orc_init_particles(self.s))

    def close(self):
        if getattr(self, "s", None):
            self.L.orc_destroy(self.s)
            self.s = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise OracleError(rc, self.L.orc_last_error(self.s).decode())

    @property
    def n_part(self):
        return self.L.orc_n_part(self.s)

    @property
    def n_ghost(self):
        return self.L.orc_n_ghost(self.s)

    @property
    def n_alive(self):
        return self.n_part - self.n_ghost

    def set_state(self, fields, n_ghost=0):
        f = np.ascontiguousarray(fields, dtype=np.float64)
        self._check(self.L.orc_set_state(self.s, f.shape[0] - n_ghost, n_ghost, f.ctypes.data))

    def fields(self):
        n = self.n_part
        f = np.zeros((n, 9), dtype=np.float64)
        self.L.orc_get_state(self.s, f.ctypes.data, None)
        return f

    def col(self, name):
        return self.fields()[:, FIELDS.index(name)]

    def step(self, nsteps=1):
        self._check(self.L.orc_step(self.s, nsteps))

    def ghosts(self):
        self._check(self.L.orc_ghosts(self.s))

    def density(self, i0=0, i1=None):
        self._check(self.L.orc_density(self.s, i0, self.n_part if i1 is None else i1))

    def forces(self, i0=0, i1=None):
        self._check(self.L.orc_forces(self.s, i0, self.n_part if i1 is None else i1))

    def kick_drift(self):
        self._check(self.L.orc_kick_drift(self.s))

    def omega(self):
        out = np.zeros(self.n_part, dtype=np.float64)
        self._check(self.L.orc_omega(self.s, out.ctypes.data))
        return out

    def newton_iters(self):
        out = np.zeros(self.n_part, dtype=np.int32)
        self.L.orc_get_newton_iters(self.s, out.ctypes.data)
        return out

    def neighbour_counts(self):
        out = np.zeros(self.n_part, dtype=np.int32)
        self.L.orc_get_neighbour_counts(self.s, out.ctypes.data)
        return out

    def fallback_count(self):
        return int(self.L.orc_fallback_count(self.s))

    def mean_force_neighbours(self):
        return float(self.L.orc_mean_force_neighbours(self.s))
