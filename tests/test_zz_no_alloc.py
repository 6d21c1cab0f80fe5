"""sphb_density_static — the one entry point of libsphb.so that allocates nothing.

Who needs it: a host program that gave the whole device to its own arena allocator at start-up
(any later cudaMalloc fails there), or one that needs a density pass inside a section where
cudaMalloc's device-wide synchronisation would stall its other streams.  Such a caller cannot
create an sphb handle (sphb_create allocates), so this entry point takes the caller's device
pointers and keeps its work arrays in a fixed array inside the library.

The test calls it through the C-ABI twice in a row, on two different particle sets of
different size, so that anything the first call left in the shared array would show in the
second, and compares h and density with the handle-based sphb_density on the same inputs.  Both
run the same kernel in the same summation order, so they must agree to the last bit; the
assertion uses the parity tolerance of the other GPU tests (1e-10 relative).  Because that is CUDA
against CUDA, the same calls are also held against the CPU oracle (oracle/, the checker), and
the two ends of the size range are covered: one particle, and exactly
SPHB_STATIC_MAX_PARTICLES, where the work arrays fill the fixed array to its last byte.  The
file is named to be collected last: a run that stops at the first failure has then run every
other test.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-10


def _inputs(n, seed):
    """A perturbed lattice: ascending x, equal masses, smoothing lengths off their roots."""
    rng = np.random.default_rng(seed)
    spacing = 4.0 / n
    x = -2.0 + spacing * (np.arange(n) + 0.5) + rng.uniform(-0.2, 0.2, n) * spacing
    assert np.all(np.diff(x) > 0)
    h = 2.0 * spacing * rng.uniform(0.9, 1.15, n)
    rho = np.ones(n)
    return x, h, rho, spacing   # mass = spacing gives density ~ 1


def _reference_entry_point(x, h, rho, mass):
    """The existing route: a handle, host records up, sphb_density, records down."""
    from sph_b200 import PARTICLE_DTYPE, Config, Simulation
    n = len(x)
    rec = np.zeros(n, dtype=PARTICLE_DTYPE)
    rec["id"] = np.arange(n)
    rec["mass"], rec["pos"], rec["h"], rec["density"] = mass, x, h, rho
    sim = Simulation(Config(n_part=n, mass=mass, pressure_calc=1))
    sim.upload(rec, 0)
    sim.density(include_ghosts=False)
    out, ng = sim.download()
    sim.close()
    assert ng == 0
    return out["h"].copy(), out["density"].copy()


def _static_entry_point(x, h, rho, mass):
    import torch
    from sph_b200 import api
    dev = torch.device("cuda", 0)
    dx, dh, drho = (torch.from_numpy(a).to(dev) for a in (x, h, rho))
    h_out = torch.full_like(dh, float("nan"))
    rho_out = torch.full_like(dh, float("nan"))
    torch.cuda.synchronize()
    api.density_static(dx.data_ptr(), dh.data_ptr(), drho.data_ptr(), len(x), mass, 2.0, 2.0, True,
                       h_out.data_ptr(), rho_out.data_ptr(), stream=0)
    return h_out.cpu().numpy(), rho_out.cpu().numpy()


def _oracle(x, h, rho, mass, variable_h=True):
    """The CPU checker on the same inputs: set_state + DensityCalculator for every particle."""
    from oracle.sph_oracle import Oracle
    n = len(x)
    f = np.zeros((n, 9))
    f[:, 0], f[:, 1], f[:, 4], f[:, 7] = mass, x, h, rho
    o = Oracle(n, mass, 1, limit=2.0, h_factor=2.0, variable_h=variable_h, init=False)
    o.set_state(f, 0)
    o.density()
    out = o.fields()
    o.close()
    return out[:, 4].copy(), out[:, 7].copy()


def _rel(a, b):
    return float(np.max(np.abs(a - b) / np.abs(b)))


def test_static_density_matches_handle_density_on_two_calls_in_a_row():
    # second call: smaller and different, so stale contents of the shared array would matter
    for n, seed in ((200_000, 7), (60_001, 8)):
        x, h, rho, mass = _inputs(n, seed)
        h_ref, rho_ref = _reference_entry_point(x, h, rho, mass)
        h_st, rho_st = _static_entry_point(x, h, rho, mass)
        assert np.isfinite(h_st).all() and np.isfinite(rho_st).all()
        assert _rel(h_st, h_ref) < TOL, "h, n=%d" % n
        assert _rel(rho_st, rho_ref) < TOL, "density, n=%d" % n
        assert (h_st != h).any()          # the root-find really ran
        h_or, rho_or = _oracle(x, h, rho, mass)
        assert _rel(h_st, h_or) < TOL, "h vs oracle, n=%d" % n
        assert _rel(rho_st, rho_or) < TOL, "density vs oracle, n=%d" % n


def test_static_density_at_the_full_size_of_its_array_and_then_a_small_call():
    """n = SPHB_STATIC_MAX_PARTICLES uses every byte of the work arrays; the small call right after it
    finds them full of the big call's leftovers."""
    from sph_b200 import api
    for n, seed in ((api.STATIC_MAX_PARTICLES, 11), (1000, 12)):
        x, h, rho, mass = _inputs(n, seed)
        h_st, rho_st = _static_entry_point(x, h, rho, mass)
        h_or, rho_or = _oracle(x, h, rho, mass)
        assert _rel(h_st, h_or) < TOL, "h vs oracle, n=%d" % n
        assert _rel(rho_st, rho_or) < TOL, "density vs oracle, n=%d" % n


def test_static_density_with_one_particle():
    """Fixed h: the density of a lone particle is m w(0) / h.  Variable h: f(h) has no root for a lone
    particle, Newton runs out of iterations and the bisection bracket does not straddle zero: GSL aborts the
    reference there (smoothing_length.cpp:165-185); the oracle and the library report it as an error."""
    import torch
    from oracle.sph_oracle import OracleError
    from sph_b200 import SphbError, api
    x, h, rho = np.array([0.25]), np.array([0.1]), np.array([1.0])
    mass = 0.04
    dev = torch.device("cuda", 0)
    dx, dh, drho = (torch.from_numpy(a).to(dev) for a in (x, h, rho))
    h_out, rho_out = torch.zeros_like(dh), torch.zeros_like(dh)
    api.density_static(dx.data_ptr(), dh.data_ptr(), drho.data_ptr(), 1, mass, 2.0, 2.0, False,
                       h_out.data_ptr(), rho_out.data_ptr())
    h_or, rho_or = _oracle(x, h, rho, mass, variable_h=False)
    assert h_out.item() == 0.1 and abs(rho_out.item() - rho_or[0]) < TOL * rho_or[0]
    assert abs(rho_or[0] - mass * (115.0 / 192.0) / 0.1) < 1e-14
    with pytest.raises(OracleError) as eo:
        _oracle(x, h, rho, mass, variable_h=True)
    assert eo.value.code == 6                      # ORC_ERR_NO_BRACKET
    with pytest.raises(SphbError) as e:
        api.density_static(dx.data_ptr(), dh.data_ptr(), drho.data_ptr(), 1, mass, 2.0, 2.0, True,
                           h_out.data_ptr(), rho_out.data_ptr())
    assert e.value.code == 9 and "straddle" in e.value.msg   # SPHB_ERR_NO_BRACKET


def test_static_density_refuses_more_particles_than_its_array_holds():
    import torch
    from sph_b200 import SphbError, api
    small = torch.zeros(16, dtype=torch.float64, device="cuda:0")
    with pytest.raises(SphbError) as e:
        api.density_static(small.data_ptr(), small.data_ptr(), small.data_ptr(), api.STATIC_MAX_PARTICLES + 1,
                           1.0, 2.0, 2.0, True, small.data_ptr(), small.data_ptr())
    assert e.value.code == 3 and "SPHB_STATIC_MAX_PARTICLES" in e.value.msg
