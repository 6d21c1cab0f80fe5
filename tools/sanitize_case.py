#!/usr/bin/env python
"""Small-N cases for compute-sanitizer (tools/sanitize.sh): the shipped configuration in both sweep strategies,
a narrow-group serial sweep that hands work to the refinement, a fixed-h run, a 2-slab run in one process, the
handle-less density entry point.  Prints one line per case; exits non-zero if a result check fails."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sph_b200 import Config, Simulation  # noqa: E402
from sph_b200.slabs import make_local_slabs  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "all"
ok = True


def shipped(mode, steps=20, **kw):
    s = Simulation(Config(n_part=101, mass=0.04, pressure_calc=1, t_i=0.005), sweep_mode=mode, **kw)
    s.init_lattice()
    s.step(steps)
    f = s.fields()[0]
    s.close()
    return f


if which in ("all", "single"):
    a = shipped(1)
    b = shipped(2)
    c = shipped(2, sweep_group=8, sweep_segment=16, sweep_burn_in=8)
    same = np.array_equal(a, b) and np.array_equal(a, c)
    print("N=101 adiabatic variable-h, 20 steps: work lists / serial sweep / short segments identical:", same)
    ok &= same
    s = Simulation(Config(n_part=401, mass=0.01, pressure_calc=0, t_i=0.00125), variable_h=False, fixed_h=0.05, sweep_mode=2)
    s.init_lattice()
    s.step(8)
    print("N=401 isothermal fixed-h, 8 steps, serial sweep G=16: n_ghost", s.fields()[2])
    s.close()
if which in ("all", "slabs"):
    n = 4001
    cfg = Config(n_part=n, mass=4.0 / (n - 1), pressure_calc=1, t_i=0.5 / (n - 1))
    one = Simulation(cfg)
    one.init_lattice()
    many = make_local_slabs(cfg, 2, halo_count=64)
    many.init_lattice()
    one.step(4)
    many.step(4)
    f1 = one.fields()[0]
    f2 = many.fields()[0]
    same = np.array_equal(f1[:n], f2[:n])
    print("N=4001 as 2 slabs in one process, 4 steps: identical to the single handle:", same)
    ok &= same
    one.close()
    many.close()
sys.exit(0 if ok else 1)
