"""x-slab domain decomposition across the GPUs of one node."""
from __future__ import annotations

from .api import Config, Simulation


def make_simulation(config: Config, variable_h=True, fixed_h=0.2, device=-1, rank=0, world=1, **kw):
    """world == 1: one device-resident Simulation.  world > 1: one slab per rank."""
    if world == 1:
        return Simulation(config, variable_h=variable_h, fixed_h=fixed_h, device=device, **kw)
    raise NotImplementedError("x-slab decomposition is not built yet")
