"""The drop-in route end to end: the reference's unchanged main.cpp / setup.cpp / kernel.cpp linked
with the C++ adaptors (sph_b200/adaptors) and libsphb.so, run as a program on the shipped
configuration, against the reference's own binary (oracle/_ref/sph_ref, the checker) — every dump
file of the 200-step run, compared at the precision the reference prints (sph_simulation.cpp:131-133).
Both binaries are built in the build container (they need the reference's sources) and travel to
the GPU box; the test is skipped where they are absent."""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DROPIN = os.path.join(ROOT, "sph_b200", "adaptors", "_build")
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "sph_ref")

# the reference's shipped config.txt (keys of setup.cpp:27-33)
CONFIG = """n_part 101
mass 0.04
pressure_calc {pc}
limit 2
v_0 1
h_factor 2
t_i 0.005
"""


def _run(binary, workdir, pc, extra=""):
    os.makedirs(workdir, exist_ok=True)
    with open(os.path.join(workdir, "config.txt"), "w") as f:
        f.write(CONFIG.format(pc=pc) + extra)
    res = subprocess.run([binary, "config.txt"], cwd=workdir, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                         text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:]
    return res.stdout


def _load(path):
    ids, types, rows = [], [], []
    for line in open(path):
        if line.startswith("#"):
            continue
        f = line.split()
        ids.append(int(f[0]))
        types.append(f[1])
        rows.append([float(v) for v in f[2:]])
    return ids, types, np.array(rows)


# Optional keys the adaptors read from the same config file (sph_b200/adaptors/sphb_bridge.hpp).  The reference's reader
# looks up only its own seven keys (setup.cpp:27-33), so BOTH programs are given the extended file: `variable_h 0` makes
# the one drop-in binary behave like the reference's fixed-h BUILD (its compile-time switch, define.hpp:27-29).
FIXED = "variable_h 0\nfixed_h 0.2\n# sweep strategy: same results either way\nsweep_mode serial\n"


@pytest.mark.parametrize("binary,ref,pc,extra", [("sph_dropin", "sph_ref", 1, ""), ("sph_dropin", "sph_ref", 0, ""),
                                                 ("sph_dropin", "sph_ref_fixedh", 1, FIXED),
                                                 ("sph_dropin", "sph_ref_fixedh", 0, FIXED)])
def test_dropin_program_reproduces_the_reference_dumps(tmp_path, binary, ref, pc, extra):
    exe = os.path.join(DROPIN, binary)
    ref_exe = os.path.join(ROOT, "oracle", "_ref", ref)
    if not (os.path.exists(exe) and os.path.exists(ref_exe)):
        pytest.skip("drop-in / reference binaries not built (need the reference sources)")
    _run(exe, str(tmp_path / "ours"), pc, extra)
    _run(ref_exe, str(tmp_path / "ref"), pc, extra)
    ours = sorted(os.listdir(tmp_path / "ours" / "dumps"), key=lambda s: int(s[:-4]))
    theirs = sorted(os.listdir(tmp_path / "ref" / "dumps"), key=lambda s: int(s[:-4]))
    assert ours == theirs and len(ours) == 201
    # printed with 5 (h) and 3 decimals: allow one unit in the last printed place (a value that
    # sits on a rounding boundary may print differently); u is never initialised by the reference in
    # isothermal runs (setup.cpp:171-172), so it is not compared there
    tol = np.array([1.01e-5] + [1.01e-3] * 6)
    ncols = 7 if pc == 1 else 6
    worst = 0.0
    for name in ours:
        ia, ta, a = _load(tmp_path / "ours" / "dumps" / name)
        ib, tb, b = _load(tmp_path / "ref" / "dumps" / name)
        assert ta == tb, "particle types / ghost count differ in %s" % name
        assert a.shape == b.shape
        # acceleration is printed with 3 decimals but reaches O(100): compare relative there
        d = np.abs(a - b)[:, :ncols]
        lim = np.maximum(tol[:ncols], 1e-9 * np.abs(b[:, :ncols]))
        assert (d <= lim).all(), "%s: %s" % (name, np.argwhere(d > lim)[:5])
        worst = max(worst, float(d.max()))
    print("worst printed difference over 201 dumps: %.2e" % worst)


def test_dropin_program_thinned_out_dumps(tmp_path):
    """`dump_every 50` + `dump_stride 10`: only steps 0, 50, ..., 200 are dumped (the state stays on the device in
    between), each with every 10th alive particle and all ghosts — the same rows as in the reference's full dumps."""
    exe = os.path.join(DROPIN, "sph_dropin")
    if not (os.path.exists(exe) and os.path.exists(REF_BIN)):
        pytest.skip("drop-in / reference binaries not built (need the reference sources)")
    _run(exe, str(tmp_path / "ours"), 1, "dump_every 50\ndump_stride 10\n")
    _run(REF_BIN, str(tmp_path / "ref"), 1)
    ours = sorted(os.listdir(tmp_path / "ours" / "dumps"), key=lambda s: int(s[:-4]))
    assert ours == ["0.txt", "50.txt", "100.txt", "150.txt", "200.txt"]
    tol = np.array([1.01e-5] + [1.01e-3] * 6)
    for name in ours:
        ia, ta, a = _load(tmp_path / "ours" / "dumps" / name)
        ib, tb, b = _load(tmp_path / "ref" / "dumps" / name)
        if name == "0.txt":
            # written by the reference's own init path before any step: the full array
            assert len(ia) == len(ib)
            continue
        alive = [k for k, t in enumerate(tb) if t == "Alive"]
        ghosts = [k for k, t in enumerate(tb) if t != "Alive"]
        rows = alive[::10] + ghosts
        assert len(ia) == len(rows) and ta == [tb[k] for k in rows]
        assert ia[:len(alive[::10])] == [ib[k] for k in alive[::10]]      # alive ids; ghost ids are allocation artefacts
        d = np.abs(a - b[rows])
        lim = np.maximum(tol, 1e-9 * np.abs(b[rows]))
        assert (d <= lim).all(), name


def test_adaptors_rethrow_with_the_reference_exception_types():
    """calculators.cpp:184 throws a POINTER to std::logic_error for a density below epsilon, ghost_particles.cpp:39
    throws std::logic_error by value for a zero smoothing length: a host program written against the reference
    catches exactly those.  sph_b200/adaptors/throw_probe.cpp provokes both through the adaptors' functors."""
    exe = os.path.join(DROPIN, "throw_probe")
    if not os.path.exists(exe):
        pytest.skip("probe not built (needs the reference sources)")
    res = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert res.returncode == 0, res.stdout
    assert "pointer: Particle had density less than epsilon!" in res.stdout
    assert "value: Ghost particle initialization: Particle had zero smoothing length!" in res.stdout
