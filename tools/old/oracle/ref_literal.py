"""oracle/ref_literal.py — TEST INFRASTRUCTURE, not product code.

ctypes view of oracle/_ref/libref_*.so: the reference's OWN sources
(/root/reference/sph/*.cpp, compiled unmodified by oracle/Makefile) behind the
small C driver in oracle/ref_harness.cpp.  Used to pin the plain-C restatement
(oracle/sph_oracle.c), to generate tests/golden/, and as the "reference" CPU
baseline of bench.py.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
FIELDS = ("mass", "pos", "vel", "acc", "h", "du_dt", "u", "density", "pressure")


def lib_path(variable_h: bool = True, flavour: str = "strict") -> str:
    name = "libref_%s%s.so" % ("" if variable_h else "fixedh_", flavour)
    return os.path.join(_HERE, "_ref", name)


def available(variable_h: bool = True, flavour: str = "strict") -> bool:
    return os.path.exists(lib_path(variable_h, flavour))


_libs: dict = {}


def _load(variable_h: bool, flavour: str):
    key = (variable_h, flavour)
    if key in _libs:
        return _libs[key]
    lib = C.CDLL(lib_path(variable_h, flavour))
    d, i, p = C.c_double, C.c_int, C.c_void_p
    lib.ref_create.restype = p
    lib.ref_create.argtypes = [i, d, i, d, d, d, d]
    lib.ref_create_raw.restype = p
    lib.ref_create_raw.argtypes = [i, i, d, i, d, d, d, d, p, p]
    lib.ref_destroy.argtypes = [p]
    lib.ref_last_error.restype = C.c_char_p
    lib.ref_last_error.argtypes = [p]
    for f in ("ref_n_part", "ref_n_ghost"):
        getattr(lib, f).argtypes = [p]
        getattr(lib, f).restype = i
    lib.ref_step.argtypes = [p, i]
    lib.ref_get_state.argtypes = [p, p, p, p]
    lib.ref_op_ghosts.argtypes = [p]
    lib.ref_op_density.argtypes = [p, i, i]
    lib.ref_op_forces.argtypes = [p, i, i]
    lib.ref_op_kick_drift.argtypes = [p]
    lib.ref_op_omega.argtypes = [p, p]
    lib.ref_kernel.restype = d
    lib.ref_kernel.argtypes = [d]
    lib.ref_dkernel_dq.restype = d
    lib.ref_dkernel_dq.argtypes = [d]
    lib.ref_newton_log.argtypes = [p, i]
    lib.ref_set_constant_h.argtypes = [d]
    _libs[key] = lib
    return lib


@dataclass
class State:
    """Array-index-ordered snapshot: alive block then ghost block."""
    n_part: int
    n_ghost: int
    fields: np.ndarray  # [n_part, 9] float64, columns = FIELDS
    type: np.ndarray    # [n_part] int32 (0 alive, 1 ghost)

    def __getattr__(self, name):
        if name in FIELDS:
            return self.fields[:, FIELDS.index(name)]
        raise AttributeError(name)

    @property
    def n_alive(self):
        return self.n_part - self.n_ghost


class RefLiteral:
    """One reference simulation (init_particles + step_forward)."""

    def __init__(self, n_part, mass, pressure_calc, limit=2.0, v_0=1.0, h_factor=2.0,
                 t_i=0.005, variable_h=True, constant_h=0.2, flavour="strict", raw=None):
        self.lib = _load(variable_h, flavour)
        if not variable_h:
            self.lib.ref_set_constant_h(constant_h)
        if raw is None:
            self.h = self.lib.ref_create(n_part, mass, pressure_calc, limit, v_0, h_factor, t_i)
        else:
            fields, ptype, n_ghost = raw
            fields = np.ascontiguousarray(fields, dtype=np.float64)
            ptype = np.ascontiguousarray(ptype, dtype=np.int32)
            self._keep = (fields, ptype)
            self.h = self.lib.ref_create_raw(n_part, n_ghost, mass, pressure_calc, limit, v_0,
                                             h_factor, t_i, fields.ctypes.data, ptype.ctypes.data)
        self._drain_log()

    def close(self):
        if self.h:
            self.lib.ref_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise RuntimeError(self.lib.ref_last_error(self.h).decode())

    def _drain_log(self):
        cap = 1 << 22
        buf = np.zeros(cap, dtype=np.int32)
        n = self.lib.ref_newton_log(buf.ctypes.data, cap)
        return buf[: min(n, cap)].copy()

    def step(self, nsteps=1):
        """Returns the Newton iteration counts of the root-finds performed, in call order."""
        self._drain_log()
        self._check(self.lib.ref_step(self.h, nsteps))
        return self._drain_log()

    def state(self) -> State:
        n = self.lib.ref_n_part(self.h)
        f = np.zeros((n, 9), dtype=np.float64)
        t = np.zeros(n, dtype=np.int32)
        self.lib.ref_get_state(self.h, f.ctypes.data, t.ctypes.data, None)
        return State(n, self.lib.ref_n_ghost(self.h), f, t)

    # function-level
    def op_ghosts(self):
        self._check(self.lib.ref_op_ghosts(self.h))

    def op_density(self, i0=0, i1=None):
        self._drain_log()
        if i1 is None:
            i1 = self.lib.ref_n_part(self.h)
        self._check(self.lib.ref_op_density(self.h, i0, i1))
        return self._drain_log()

    def op_forces(self, i0=0, i1=None):
        if i1 is None:
            i1 = self.lib.ref_n_part(self.h)
        self._check(self.lib.ref_op_forces(self.h, i0, i1))

    def op_kick_drift(self):
        self._check(self.lib.ref_op_kick_drift(self.h))

    def op_omega(self):
        n = self.lib.ref_n_part(self.h)
        out = np.zeros(n, dtype=np.float64)
        self._check(self.lib.ref_op_omega(self.h, out.ctypes.data))
        return out


def kernel_w(q, flavour="strict"):
    return _load(True, flavour).ref_kernel(float(q))


def kernel_dw(q, flavour="strict"):
    return _load(True, flavour).ref_dkernel_dq(float(q))
