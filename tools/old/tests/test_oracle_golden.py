"""CPU tests: the plain-C oracle (oracle/sph_oracle.c) against the fixtures that the
reference's own code produced (tests/golden/, see make_golden.py), against the survey's
fingerprints and against hand-computed kernel values."""
import numpy as np
import pytest

from conftest import FIELDS, golden_kwargs, golden_names, load_golden, scaled_err
from oracle.sph_oracle import Oracle, lib


def _run(name, exact):
    g = load_golden(name)
    o = Oracle(index_order_sums=exact, **golden_kwargs(g))
    steps = [int(s) for s in g["steps"]]
    timeline = g["n_ghost_timeline"]
    for st in range(0, max(steps) + 1):
        if st > 0:
            o.step(1)
        assert o.n_ghost == int(timeline[st]), "n_ghost at step %d" % st
        if st in steps:
            yield st, o, g["fields_%d" % st], g["type_%d" % st], (g["newton_%d" % st] if st > 0 else None)


@pytest.mark.parametrize("name", golden_names())
def test_index_order_mode_is_bit_identical_to_reference(name):
    """In ascending-index summation order the restatement reproduces the reference's strict
    build bit for bit: every field of every alive and ghost particle, ghost counts at every
    step, Newton iteration counts."""
    for st, o, ref, typ, newton in _run(name, exact=True):
        f = o.fields()
        assert f.shape == ref.shape
        assert np.array_equal(f, ref), "step %d differs" % st
        na = o.n_alive
        assert np.array_equal(typ, (np.arange(f.shape[0]) >= na).astype(np.int32))
        if newton is not None and o.L.orc_fallback_count(o.s) == 0:
            got = o.newton_iters()[:na]
            if int(load_golden(name)["variable_h"]):
                assert np.array_equal(got, newton.astype(np.int32))


@pytest.mark.parametrize("name", golden_names())
def test_x_order_mode_within_tolerance(name):
    """x-ordered sums (the fast mode used as CPU baseline and at large N): rho, h, v, u and
    everything else within 1e-10 of the reference, counts equal."""
    for st, o, ref, typ, newton in _run(name, exact=False):
        f = o.fields()
        for c, nm in enumerate(FIELDS):
            assert scaled_err(f[:, c], ref[:, c]) < 1e-10, (nm, st)


def test_survey_fingerprints_shipped_config():
    """SURVEY.md Appendix C, strict build of the reference on the shipped config.txt."""
    o = Oracle(101, 0.04, 1, index_order_sums=True)
    f = o.fields()
    assert o.n_ghost == 10 and o.n_part == 111
    col = {n: f[:, i] for i, n in enumerate(FIELDS)}
    assert col["pos"][0] == -1.98
    assert col["vel"][0] == 1.2909944487358056
    assert col["h"][0] == 0.079428910271459291
    assert col["density"][0] == 1.0071899417447034
    assert col["acc"][0] == 1.6721580022197775
    assert col["du_dt"][0] == -19.432524441326599
    assert col["pos"][50] == -2.2204460492503131e-16
    assert col["acc"][50] == -232.41609196882015
    o.step(1)
    f = o.fields()
    assert f[50, 1] == 0.0035497710940685539 and f[50, 2] == 0.44344693994631978
    assert f[50, 4] == 0.072787001504491772 and f[50, 7] == 1.0990973392268586
    assert f[0, 2] == 1.29052403689971 and f[100, 2] == -1.2910445580302967
    o.step(199)
    f = o.fields()
    assert o.n_ghost == 8
    assert f[0, 1] == -1.8460023227049884 and f[0, 2] == 0.039452420841717815
    assert f[0, 4] == 0.26962429727796011 and f[0, 7] == 0.29670916455443702
    assert f[50, 4] == 0.058326590660432935 and f[50, 6] == 3.806913804919037


def test_kernel_known_answers():
    """M5 quartic spline, reference sph/kernel.cpp:14-64: hand-computed values at the
    branch points (the reference's own unit test still holds M4 cubic numbers)."""
    L = lib()
    assert L.orc_w(0.0) == pytest.approx(115.0 / 192.0, rel=1e-15)
    assert L.orc_w(0.5) == pytest.approx((2.0 ** 4 - 5.0) / 24.0, rel=1e-15)
    assert L.orc_w(1.5) == pytest.approx(1.0 / 24.0, rel=1e-15)
    assert L.orc_w(2.5) == 0.0 and L.orc_w(3.0) == 0.0
    assert L.orc_dw(0.0) == 0.0
    assert L.orc_dw(0.5) == pytest.approx((-4 * 8.0 + 20.0) / 24.0, rel=1e-15)
    assert L.orc_dw(1.5) == pytest.approx(-4.0 / 24.0, rel=1e-15)
    assert L.orc_dw(2.5) == 0.0
    # normalisation: integral of w over [-2.5, 2.5] is 1
    q = np.linspace(-2.5, 2.5, 200001)
    w = np.array([L.orc_w(abs(x)) for x in q])
    assert np.trapezoid(w, q) == pytest.approx(1.0, abs=1e-9)


def test_three_particle_density_hand_sum():
    """x = -1/2, 0, 1/2, m = 1, fixed h = 1 (the layout of reference
    unittest/test_calculator.cpp:24-28) with the SHIPPED M5 kernel:
    rho_mid = w(0) + 2 w(1/2), rho_edge = w(0) + w(1/2) + w(1)."""
    o = Oracle(3, 1.0, 0, limit=1.0, variable_h=False, constant_h=1.0, init=False)
    f = np.zeros((3, 9))
    f[:, 0] = 1.0
    f[:, 1] = [-0.5, 0.0, 0.5]
    f[:, 4] = 1.0
    o.set_state(f)
    o.density()
    w0, wh, w1 = 115.0 / 192.0, 11.0 / 24.0, (1.5 ** 4 - 5 * 0.5 ** 4) / 24.0
    rho = o.col("density")
    assert rho[1] == pytest.approx(w0 + 2 * wh, rel=1e-15)
    assert rho[0] == pytest.approx(w0 + wh + w1, rel=1e-15)
    assert rho[2] == pytest.approx(w0 + wh + w1, rel=1e-15)
    assert list(o.neighbour_counts()) == [3, 3, 3]


def test_jacobi_order_differs_from_reference_order():
    """a compute-all-then-update-all step is NOT the reference."""
    a = Oracle(101, 0.04, 1)
    b = Oracle(101, 0.04, 1, order=1)
    a.step(1)
    b.step(1)
    fa, fb = a.fields(), b.fields()
    assert np.array_equal(fa[:, 1], fb[:, 1])            # positions agree (drift only)
    assert scaled_err(fa[:101, 4], fb[:101, 4]) < 1e-14  # h, rho do not depend on the order (F3)
    assert scaled_err(fa[:101, 7], fb[:101, 7]) < 1e-14
    assert scaled_err(fa[:101, 2], fb[:101, 2]) > 1e-3   # v does


def test_momentum_drift_is_small():
    """The check the reference's author wished for (readme.md:23).  Each stream carries
    |p| ~ 2.6; the mirror walls and the ordered sweep (F2) exchange a little momentum, so the
    total is only approximately conserved."""
    o = Oracle(101, 0.04, 1)
    f = o.fields()
    p0 = (f[:101, 0] * f[:101, 2]).sum()
    o.step(50)
    f = o.fields()
    p1 = (f[:101, 0] * f[:101, 2]).sum()
    assert abs(p1 - p0) < 0.1
