"""Error paths of the particle loop: what the reference throws (calculators.cpp:38-42, 174-186,
ghost_particles.cpp:37-43) and where GSL's default handler aborts it (smoothing_length.cpp:165-185, 232-257;
SURVEY.md Appendix B) must come back from the device as the matching status code, with the reference's message
and the FIRST failing particle — the one the reference's sequential loops reach first.  The oracle (the plain-C
restatement, which returns codes where the reference throws / aborts) is driven on the same inputs.

Code 4 (negative smoothing length, calculators.cpp:38-42) has no test: the guard cannot fire.  rootfind_h returns
either a Newton iterate that passed gsl_root_test_delta or the bisection root inside [1e-8, 2 limit]; a Newton
start is replaced by the positive default when the stored h is not above 1e-8, and from a positive start the
iteration f(h) = rho_sum(h) - m hf / h either converges upwards or diverges to the bisection.  A search over 4000
random small states with the oracle (2-7 particles, h over four decades, five h_factors) found none.  The
device carries the check all the same.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

# library code -> (oracle code, message fragment)
LOW_DENSITY, ZERO_DERIVATIVE, NOT_FINITE, NO_BRACKET = 5, 7, 8, 9
ORC = {LOW_DENSITY: 2, ZERO_DERIVATIVE: 4, NOT_FINITE: 5, NO_BRACKET: 6}


def _records(x, h, mass, rho=1.0, u=1.5):
    from sph_b200 import PARTICLE_DTYPE
    n = len(x)
    rec = np.zeros(n, dtype=PARTICLE_DTYPE)
    rec["id"] = np.arange(n)
    rec["mass"], rec["pos"], rec["h"], rec["density"], rec["u"], rec["pressure"] = mass, x, h, rho, u, 1.0
    return rec


def _both(rec, mass, op, variable_h=True, h_factor=2.0, pc=1):
    """Run `op` on the oracle and on the device from the same records; return (oracle error, library error)."""
    from oracle.sph_oracle import Oracle, OracleError
    from sph_b200 import Config, Simulation, SphbError
    from conftest import FIELDS
    n = len(rec)
    f = np.stack([rec[name] for name in FIELDS], axis=1)
    o = Oracle(n, mass, pc, limit=2.0, h_factor=h_factor, variable_h=variable_h, constant_h=0.2, init=False)
    o.set_state(f, 0)
    eo = es = None
    try:
        getattr(o, op)() if op != "step" else o.step(1)
    except OracleError as e:
        eo = e
    o.close()
    sim = Simulation(Config(n_part=n, mass=mass, pressure_calc=pc, h_factor=h_factor), variable_h=variable_h, fixed_h=0.2)
    sim.upload(rec, 0)
    try:
        getattr(sim, op)() if op != "step" else sim.step(1)
    except SphbError as e:
        es = e
    # the handle stays usable after an error has been reported
    assert sim.counters().n_alive == n
    sim.close()
    return eo, es


def _particle(msg):
    import re
    m = re.search(r"particle (\d+)", msg)
    return int(m.group(1)) if m else None


def test_zero_derivative_like_gsl_newton():
    """All masses zero: f = f' = 0, GSL's newton step reports 'derivative is zero' and aborts the reference."""
    x = np.linspace(-1.9, 1.9, 50)
    eo, es = _both(_records(x, 0.16, 0.0), 0.0, "density")
    assert eo is not None and eo.code == ORC[ZERO_DERIVATIVE]
    assert es is not None and es.code == ZERO_DERIVATIVE and "derivative is zero" in es.msg
    assert _particle(es.msg) == _particle(str(eo)) == 0


def test_not_finite_like_gsl_newton():
    """Infinite masses: f = inf - inf is not a number at the first Newton step of particle 0."""
    x = np.linspace(-1.9, 1.9, 50)
    eo, es = _both(_records(x, 0.16, np.inf), np.inf, "density")
    assert eo is not None and eo.code == ORC[NOT_FINITE]
    assert es is not None and es.code == NOT_FINITE and "not finite" in es.msg
    assert _particle(es.msg) == _particle(str(eo)) == 0


def test_no_bracket_like_gsl_bisection():
    """A negative h_factor: f > 0 everywhere, Newton runs out of iterations, and the bisection bracket
    [1e-8, 2 limit] does not straddle zero."""
    x = np.linspace(-1.9, 1.9, 50)
    eo, es = _both(_records(x, 0.16, 0.08), 0.08, "density", h_factor=-2.0)
    assert eo is not None and eo.code == ORC[NO_BRACKET]
    assert es is not None and es.code == NO_BRACKET and "straddle" in es.msg
    assert _particle(es.msg) == _particle(str(eo)) == 0


@pytest.mark.parametrize("variable_h", [True, False])
def test_low_density_like_ensure_nonzero_density(variable_h):
    """Masses of 1e-12 give densities of ~1e-11 < CALC_EPSILON: the reference throws while it processes particle 0
    (calculators.cpp:174-186 guards every particle the acceleration loop touches)."""
    x = np.linspace(-1.9, 1.9, 50)
    eo, es = _both(_records(x, 0.2 if not variable_h else 0.16, 1e-12), 1e-12, "step", variable_h=variable_h)
    assert eo is not None and eo.code == ORC[LOW_DENSITY]
    assert es is not None and es.code == LOW_DENSITY and "density less than epsilon" in es.msg
    assert _particle(es.msg) == _particle(str(eo)) == 0


def test_the_first_failing_particle_is_reported_not_a_random_one():
    """Every particle from 7 on has zero mass.  Thousands of threads raise (zero derivative from particle 12 on, where no
    massive neighbour is in reach; zero density from 7 on); the reference's loops would fail first at particle 7, whose
    root-find still works (massive neighbours) but whose density is m_i * sum = 0: that one is named, every time."""
    n = 20_000
    x = np.linspace(-1.99, 1.99, n)
    mass = np.full(n, 4.0 / n)
    mass[7:] = 0.0
    from sph_b200 import Config, Simulation, SphbError
    sim = Simulation(Config(n_part=n, mass=4.0 / n, pressure_calc=1))
    sim.upload(_records(x, 2 * 4.0 / n, mass), 0)
    for _ in range(3):
        with pytest.raises(SphbError) as es:
            sim.density()
        assert es.value.code == LOW_DENSITY and _particle(es.value.msg) == 7, es.value.msg
    sim.close()
