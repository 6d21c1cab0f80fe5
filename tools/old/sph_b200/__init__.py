"""sph_b200 — B200-native per-timestep particle loop of SapiensAnatis/sph.

The product is libsphb.so (hand-written sm_100a CUDA behind the C-ABI of include/sphb.h).
This package is the thin Python host side: a ctypes binding plus a mirror of the
reference's interface names (Config, Particle records, DensityCalculator, ...), used by the
tests and by bench.py.  There is NO CPU fallback: without the built library and a CUDA
device every call raises.
"""
from .api import (  # noqa: F401
    ADIABATIC, ISOTHERMAL, ORDER_JACOBI, ORDER_REFERENCE, PARTICLE_DTYPE, Config, Counters, SphbError,
    Simulation, exported_symbols, lib_path, load_library,
)
from .reference_api import (  # noqa: F401
    AccelerationCalculator, DensityCalculator, EnergyCalculator, SPHSimulation, init_particles,
    setup_ghost_particles,
)
