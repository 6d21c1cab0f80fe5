"""Python mirror of the reference's operator interface for the hot path, on top of the C-ABI.

Same names, argument meaning and error behaviour as the reference so that the parity tests
read like the reference's own tests:

    reference (C++)                                         here
    ------------------------------------------------------  ---------------------------------
    Config, Particle[]  (sph/basictypes.hpp)                 Config, numpy array of PARTICLE_DTYPE
    DensityCalculator(config, p_arr); dc(p); dc.update(..)   DensityCalculator(config, p_arr); dc(i)
    AccelerationCalculator / EnergyCalculator                idem; ac.sound_speed(i)
    setup_ghost_particles(p_arr, config)                     returns (p_arr, config), both replaced
    init_particles(config, p_arr)                            returns (p_arr, config)
    SPHSimulation(config, p_arr).step_forward()              SPHSimulation(config, p_arr).step_forward()

A functor call works on ONE particle of a host array the caller may have edited at will, so
each call uploads the array, runs the batch kernel and writes that particle's results back
(the reference pays O(N) per call too).  `SPHSimulation` is the efficient route: the state
stays in HBM across steps.  Errors the reference throws as std::logic_error arrive as
SphbError with the reference's message.  The C++ adaptors with the same names live in
sph_b200/adaptors/.
"""
from __future__ import annotations

import copy

import numpy as np

from .api import ADIABATIC, ISOTHERMAL, PARTICLE_DTYPE, Config, Simulation

GAMMA = 5.0 / 3.0  # sph/define.hpp:32


class Calculator:
    """sph/calculators.hpp:19-44."""

    def __init__(self, config: Config, p_arr: np.ndarray, variable_h=True):
        self.variable_h = variable_h
        self.update(config, p_arr)

    def update(self, config: Config, p_arr: np.ndarray):
        """Calculator::update (calculators.hpp:31-34): new Config copy + array."""
        assert p_arr.dtype == PARTICLE_DTYPE
        self.config = copy.copy(config)
        self.p_arr = p_arr

    def _sim(self) -> Simulation:
        c = self.config
        sim = Simulation(c, variable_h=self.variable_h)
        sim.upload(np.ascontiguousarray(self.p_arr[: c.n_part]), c.n_ghost)
        return sim

    def __call__(self, i: int):
        raise NotImplementedError("Attempt to call un-implemented operator() function!")


class DensityCalculator(Calculator):
    """sph/calculators.cpp:33-55: p.h <- rootfind_h(p); p.density <- calc_density(p, p.h)."""

    def __call__(self, i: int):
        sim = self._sim()
        try:
            sim.density(include_ghosts=True)
            out, _ = sim.download()
            self.p_arr["h"][i] = out["h"][i]
            self.p_arr["density"][i] = out["density"][i]
        finally:
            sim.close()

    def all(self, include_ghosts=True):
        """dc(p) for every particle (order-independent)."""
        sim = self._sim()
        try:
            sim.density(include_ghosts=include_ghosts)
            out, _ = sim.download()
            n = self.config.n_part if include_ghosts else self.config.n_part - self.config.n_ghost
            self.p_arr["h"][:n] = out["h"][:n]
            self.p_arr["density"][:n] = out["density"][:n]
        finally:
            sim.close()


class AccelerationCalculator(Calculator):
    """sph/calculators.cpp:63-129: p.pressure, p.acc (ghosts are skipped)."""

    def __call__(self, i: int):
        if self.p_arr["type"][i] == 1:
            return
        sim = self._sim()
        try:
            sim.forces()
            out, _ = sim.download()
            self.p_arr["pressure"][i] = out["pressure"][i]
            self.p_arr["acc"][i] = out["acc"][i]
        finally:
            sim.close()

    def sound_speed(self, i: int) -> float:
        """sph/calculators.cpp:139-146 (host arithmetic on one record)."""
        if self.config.pressure_calc == ISOTHERMAL:
            return 1.0
        if self.config.pressure_calc == ADIABATIC:
            return float(np.sqrt(GAMMA * self.p_arr["pressure"][i] / self.p_arr["density"][i]))
        raise ValueError("Unknown pressure calculation mode!")


class EnergyCalculator(AccelerationCalculator):
    """sph/calculators.cpp:192-213: p.du_dt from the stored pressure."""

    def __call__(self, i: int):
        sim = self._sim()
        try:
            sim.forces()
            out, _ = sim.download()
            self.p_arr["du_dt"][i] = out["du_dt"][i]
        finally:
            sim.close()


def setup_ghost_particles(p_arr: np.ndarray, config: Config, variable_h=True):
    """sph/ghost_particles.cpp:19-115.  The reference mutates both arguments (it may
    reallocate the array and rewrites n_part / n_ghost); here both are returned."""
    n_alive = config.n_part - config.n_ghost
    c = copy.copy(config)
    c.n_part, c.n_ghost = n_alive, 0
    sim = Simulation(c, variable_h=variable_h)
    try:
        sim.upload(np.ascontiguousarray(p_arr[:n_alive]), 0)
        g = sim.ghosts()
        out, g = sim.download()
    finally:
        sim.close()
    new = np.zeros(n_alive + g, dtype=PARTICLE_DTYPE)
    new[:] = out
    c.n_part, c.n_ghost = n_alive + g, g
    return new, c


def init_particles(config: Config, variable_h=True, fixed_h=0.2):
    """sph/setup.cpp:143-228 on the device; returns (p_arr, config) like the reference
    leaves them (array grown by the ghosts, n_part / n_ghost updated)."""
    n_alive = config.n_part
    c = copy.copy(config)
    c.n_ghost = 0
    sim = Simulation(c, variable_h=variable_h, fixed_h=fixed_h)
    try:
        sim.init_lattice()
        out, g = sim.download()
    finally:
        sim.close()
    c.n_part, c.n_ghost = n_alive + g, g
    return out.copy(), c


class SPHSimulation:
    """sph/sph_simulation.hpp:19-51: step_forward() keeps the state in HBM; p_arr and config
    are refreshed from the device after every call, as file_write() needs them."""

    def __init__(self, config: Config, p_arr: np.ndarray, variable_h=True, order=0):
        self.config = copy.copy(config)
        self.sim = Simulation(self.config, variable_h=variable_h, order=order)
        self.sim.upload(np.ascontiguousarray(p_arr[: config.n_part]), config.n_ghost)
        self.p_arr = p_arr
        self.current_time = 0.0
        self.timestep = config.t_i

    def step_forward(self, nsteps=1):
        self.sim.step(nsteps)
        self.current_time += nsteps * self.timestep
        out, g = self.sim.download()
        self.p_arr = out.copy()
        self.config.n_ghost = g
        self.config.n_part = len(out)
        return self.p_arr

    def close(self):
        self.sim.close()
