"""Driver entry points: build() compiles everything, smoke() runs one tiny step on cuda:0."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def build() -> None:
    """Compile libsphb.so for sm_100a (nvcc cross-compiles without a GPU), the plain-C oracle
    (the checker), and — only where /root/reference exists, i.e. in the build container — the
    reference itself into oracle/_ref/.  Building the checker is not using it."""
    from sph_b200 import build as b
    b.build(force=True)
    env = dict(os.environ)
    env.pop("CXX", None)
    env.pop("CC", None)
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], env=env,
                          stdout=subprocess.DEVNULL)
    if os.path.isdir("/root/reference/sph"):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "ref"], env=env,
                              stdout=subprocess.DEVNULL)
    import sph_b200  # noqa: F401
    from sph_b200.api import declared_symbols, exported_symbols
    missing = set(declared_symbols()) - set(exported_symbols())
    if missing:
        raise RuntimeError("libsphb.so misses symbols declared in include/sphb.h: %s" % sorted(missing))


def smoke() -> None:
    """One small invocation of the hot path on cuda:0, checked against the oracle:
    the shipped config (N = 101, adiabatic, variable h), t = 0 set-up and 3 steps."""
    import numpy as np
    from sph_b200 import Config, Simulation
    from oracle.sph_oracle import Oracle

    cfg = Config(n_part=101, mass=0.04, pressure_calc=1, t_i=0.005)
    sim = Simulation(cfg, device=0)
    sim.init_lattice()
    sim.step(3)
    f, t, g = sim.fields()
    o = Oracle(101, 0.04, 1, t_i=0.005)
    o.step(3)
    ref = o.fields()
    assert g == o.n_ghost, (g, o.n_ghost)
    for c, name in ((1, "pos"), (2, "vel"), (4, "h"), (6, "u"), (7, "density")):
        a, b = f[:101, c], ref[:101, c]
        scale = np.maximum(np.abs(b), 1e-3 * np.abs(b).max())
        err = float((np.abs(a - b) / scale).max())
        assert err < 1e-10, (name, err)
    cnt = sim.counters()
    print("smoke ok: n_ghost=%d sweeps_last=%d" % (g, cnt.sweeps_last))
    sim.close()


if __name__ == "__main__":
    build()
    if "--smoke" in sys.argv:
        smoke()
