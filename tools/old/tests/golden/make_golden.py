#!/usr/bin/env python
"""Generate tests/golden/*.npz from the reference's own code.

Run in the build container only (it needs oracle/_ref/libref_*strict.so, i.e.
/root/reference compiled by `make -C oracle ref`; strict build = -O2
-ffp-contract=off).  The fixtures are small and committed, because
/root/reference does not exist on the GPU box.

    python tests/golden/make_golden.py

Each file holds one run: the Config, the full double-precision state
(array-index order, alive block then ghost block) at selected steps, the
Newton iteration count of every root-find of those steps (alive-index order),
and n_ghost after every step.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.ref_literal import RefLiteral  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def scaled(n):
    """same physical problem at resolution n."""
    return dict(mass=4.0 / (n - 1), t_i=0.005 * 100.0 / (n - 1))


CASES = {
    # name: (n_part, pressure_calc, variable_h, constant_h, mass, t_i, steps to record)
    "shipped_adiabatic_varh_n101": (101, 1, True, 0.2, 0.04, 0.005, [0, 1, 2, 3, 10, 50, 200]),
    "isothermal_varh_n101": (101, 0, True, 0.2, 0.04, 0.005, [0, 1, 2, 10, 200]),
    "adiabatic_fixedh_n101": (101, 1, False, 0.2, 0.04, 0.005, [0, 1, 2, 10, 200]),
    "isothermal_fixedh_n101": (101, 0, False, 0.2, 0.04, 0.005, [0, 1, 2, 10, 200]),
    "adiabatic_varh_n100": (100, 1, True, 0.2, 4.0 / 99, 0.005 * 100 / 99, [0, 1, 2, 10, 40]),
    "adiabatic_varh_n401": (401, 1, True, 0.2, scaled(401)["mass"], scaled(401)["t_i"], [0, 1, 2, 10, 40]),
    "isothermal_fixedh_n401": (401, 0, False, 0.2 * 100 / 400, scaled(401)["mass"], scaled(401)["t_i"], [0, 1, 2, 10, 40]),
    "adiabatic_varh_n801": (801, 1, True, 0.2, scaled(801)["mass"], scaled(801)["t_i"], [0, 1, 2, 5]),
}


def run_case(name, n, pc, varh, ch, mass, dt, steps):
    sim = RefLiteral(n, mass, pc, limit=2.0, v_0=1.0, h_factor=2.0, t_i=dt, variable_h=varh,
                     constant_h=ch, flavour="strict")
    out = dict(n_part=n, pressure_calc=pc, variable_h=int(varh), constant_h=ch, mass=mass,
               limit=2.0, v_0=1.0, h_factor=2.0, t_i=dt, steps=np.array(steps, dtype=np.int32))
    timeline = []
    last = max(steps)
    s = sim.state()
    timeline.append(s.n_ghost)
    if 0 in steps:
        out["fields_0"] = s.fields.copy()
        out["type_0"] = s.type.copy()
    for k in range(1, last + 1):
        iters = sim.step(1)
        s = sim.state()
        timeline.append(s.n_ghost)
        if k in steps:
            out["fields_%d" % k] = s.fields.copy()
            out["type_%d" % k] = s.type.copy()
            out["newton_%d" % k] = iters.astype(np.int8)
    out["n_ghost_timeline"] = np.array(timeline, dtype=np.int32)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(name, "n_ghost", timeline[0], "->", timeline[-1], "steps", steps)
    sim.close()


if __name__ == "__main__":
    only = sys.argv[1:]
    for name, args in CASES.items():
        if only and name not in only:
            continue
        run_case(name, *args)
