"""The ordered sweep (sph_simulation.cpp:94-107) has two implementations in libsphb.so: work lists
(first guess for everybody, fixed-point refinement of the tiles in reach of a changed particle) and the
serial sweep (groups of lanes walk segments of the x-sorted array left to right, one pass).  Both must
deliver the reference's sequential result, so they must agree with EACH OTHER to the last bit on every
field, whatever the segment length, the burn-in or the group width — including settings chosen to make
the serial sweep's own checks fail (a burn-in far too short, a group too narrow for the neighbourhood),
where the work-list refinement has to repair what the sweep could not settle.  Parity with the oracle is
asserted on top for the serial sweep alone.
"""
import numpy as np
import pytest

from conftest import FIELDS, scaled_err

pytestmark = pytest.mark.gpu

COLS = {n: i for i, n in enumerate(FIELDS)}
WORKLIST, SERIAL = 1, 2


def _kw(n, pc=1, varh=True):
    return dict(n_part=n, mass=4.0 / (n - 1), pressure_calc=pc, limit=2.0, v_0=1.0, h_factor=2.0,
                t_i=0.005 * 100 / (n - 1), variable_h=varh, constant_h=0.2 * 100 / (n - 1))


def _sim(kw, **extra):
    from sph_b200 import Config, Simulation
    cfg = Config(n_part=kw["n_part"], mass=kw["mass"], pressure_calc=kw["pressure_calc"],
                 limit=kw["limit"], v_0=kw["v_0"], h_factor=kw["h_factor"], t_i=kw["t_i"])
    return Simulation(cfg, variable_h=kw["variable_h"], fixed_h=kw["constant_h"], **extra)


def _run(kw, steps, **extra):
    sim = _sim(kw, **extra)
    sim.init_lattice()
    sim.step(steps)
    f, typ, ng = sim.fields()
    c = sim.counters()
    sim.close()
    return f, ng, c


@pytest.mark.parametrize("n,pc,varh,steps", [(101, 1, True, 200), (2001, 1, True, 40), (2001, 0, True, 40),
                                             (5001, 0, False, 20), (100_003, 1, True, 10)])
def test_serial_sweep_and_work_lists_agree_bitwise(n, pc, varh, steps):
    kw = _kw(n, pc, varh)
    if n == 101:
        kw.update(mass=0.04, t_i=0.005)          # the shipped config.txt
    fa, ga, ca = _run(kw, steps, sweep_mode=WORKLIST)
    fb, gb, cb = _run(kw, steps, sweep_mode=SERIAL)
    assert ca.sweep_steps == 0 and cb.sweep_steps == steps
    assert ga == gb
    for c in FIELDS:
        assert np.array_equal(fa[:, COLS[c]], fb[:, COLS[c]]), "%s differs between the two sweep strategies" % c


@pytest.mark.parametrize("group,segment,burn_in", [(8, 64, 8), (8, 256, 16), (16, 128, 16), (16, 4096, 256), (8, 8, 8)])
def test_serial_sweep_geometry_does_not_change_a_bit(group, segment, burn_in):
    """Short burn-ins make the burn-in check fail (the counters say so); narrow groups meet neighbourhoods
    wider than they can hold.  The result must not change."""
    kw = _kw(20_011)
    ref, gref, _ = _run(kw, 8, sweep_mode=WORKLIST)
    f, g, c = _run(kw, 8, sweep_mode=SERIAL, sweep_group=group, sweep_segment=segment, sweep_burn_in=burn_in)
    assert c.sweep_group == group and c.sweep_segment == segment and c.sweep_burn_in == burn_in
    assert g == gref and np.array_equal(f, ref)


def test_too_short_a_burn_in_is_caught_by_its_check():
    kw = _kw(20_011)
    sim = _sim(kw, sweep_mode=SERIAL, sweep_group=8, sweep_segment=64, sweep_burn_in=8)
    sim.init_lattice()
    sim.step(3)
    c = sim.counters()
    sim.close()
    assert c.sweep_verify_fail > 0          # 8 particles cannot forget a wrong left boundary
    sim = _sim(kw, sweep_mode=SERIAL)
    sim.init_lattice()
    sim.step(3)
    c = sim.counters()
    sim.close()
    assert c.sweep_verify_fail == 0, "the default burn-in (%d) failed its check %d times" % (c.sweep_burn_in, c.sweep_verify_fail)


def test_fixed_h_with_narrow_group_falls_back_to_the_work_lists():
    """Fixed h = 5 spacings: 12 neighbours per side, more than 8 lanes can hold: every particle is handed to the
    work-list refinement, and the result is still the same."""
    kw = _kw(5001, 0, False)
    ref, gref, _ = _run(kw, 6, sweep_mode=WORKLIST)
    f, g, c = _run(kw, 6, sweep_mode=SERIAL, sweep_group=8)
    assert c.sweep_flagged > 4000
    assert g == gref and np.array_equal(f, ref)


def test_developed_flow_both_strategies_and_the_oracle():
    """N = 2001 to t = 0.25: particles have crossed (index order != x order), h spans a factor of several."""
    from oracle.sph_oracle import Oracle
    n = 2001
    kw = _kw(n)
    fa, ga, _ = _run(kw, 1000, sweep_mode=WORKLIST)
    fb, gb, cb = _run(kw, 1000, sweep_mode=SERIAL)
    assert ga == gb and np.array_equal(fa, fb)
    o = Oracle(**kw)
    o.step(1000)
    ref = o.fields()
    assert gb == o.n_ghost
    for c in ("pos", "vel", "h", "u", "density"):
        assert scaled_err(fb[:n, COLS[c]], ref[:n, COLS[c]]) < 1e-8, c


def test_auto_mode_is_the_same_bits_and_keeps_the_serial_sweep_on_a_generic_lattice():
    """A generic N gives every particle of the undisturbed streams a rounding-level kick (the lattice positions
    are not exactly equidistant in floating point): work lists would re-sweep every tile, the serial sweep stays on."""
    kw = _kw(60_000)
    ref, gref, _ = _run(kw, 6, sweep_mode=WORKLIST)
    f, g, c = _run(kw, 6)
    assert c.sweep_steps == 6
    assert g == gref and np.array_equal(f, ref)


def test_developed_start_state_against_the_oracle_and_between_strategies():
    """sphb_init_developed: the exact two-shock / two-rarefaction state at t* = 0.5, N = 100,001.  Mass, the jump
    conditions and the symmetry of the sampled state are checked directly; then the state is handed to the oracle and
    both advance 5 steps (every particle is accelerated here: no quiet streams), in both sweep strategies."""
    from oracle.sph_oracle import Oracle
    n = 100_001
    kw = _kw(n)
    runs = {}
    for mode in (WORKLIST, SERIAL):
        sim = _sim(kw, sweep_mode=mode)
        sim.init_developed(0.5)
        f0, t0, g0 = sim.fields()
        sim.step(5)
        runs[mode] = (f0, g0, sim.fields()[0], sim.counters())
        sim.close()
    f0, g0, f5, c = runs[SERIAL]
    assert np.array_equal(runs[WORKLIST][2], f5) and np.array_equal(runs[WORKLIST][0], f0)
    x, v, rho, P, u, h = (f0[:n, COLS[k]] for k in ("pos", "vel", "density", "pressure", "u", "h"))
    assert np.all(np.diff(x) > 0) and abs(x[n // 2]) < 1e-12 and np.allclose(x, -x[::-1], atol=1e-12) and np.allclose(v, -v[::-1], atol=1e-12)
    mid = slice(n // 2 - 2000, n // 2 - 200)                     # shocked gas at rest: rho1 / rho0 = 2.1514..., P1 = 4.1142...
    assert abs(np.median(rho[mid]) / 2.151409 - 1) < 2e-3 and np.abs(v[mid]).max() < 1e-12
    assert abs(np.median((2.0 / 3.0) * u[mid] * rho[mid]) / 4.114236 - 1) < 2e-3
    wall = slice(200, 2000)                                      # gas at rest at the wall: rho_w = (2/3)^3 rho0
    assert abs(np.median(rho[wall]) / (8.0 / 27.0) - 1) < 2e-3
    o = Oracle(**kw, init=False)
    o.set_state(f0, g0)
    o.step(5)
    ref = o.fields()
    assert ref.shape == f5.shape
    for col in ("pos", "vel", "h", "u", "density"):
        assert scaled_err(f5[:n, COLS[col]], ref[:n, COLS[col]]) < 1e-10, col
    assert c.sweep_steps == 5
