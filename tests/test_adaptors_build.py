"""CPU test of the C++ side of the boundary: the adaptors compile against the reference's OWN headers and
link, together with the reference's unchanged main.cpp / setup.cpp / kernel.cpp, into the drop-in program
(only where the reference sources exist, i.e. in the build container; the GPU run of that program is
tests/test_dropin_binary.py)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ADAPTORS = os.path.join(ROOT, "sph_b200", "adaptors")


@pytest.mark.skipif(not os.path.isdir("/root/reference/sph"), reason="reference sources not present")
def test_dropin_program_links_against_reference_headers():
    import sph_b200.build as b
    b.build()
    env = dict(os.environ)
    env.pop("CXX", None)
    env.pop("CC", None)
    res = subprocess.run(["make", "-C", ADAPTORS], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert res.returncode == 0, res.stdout[-3000:]
    for exe in ("sph_dropin",):
        path = os.path.join(ADAPTORS, "_build", exe)
        assert os.path.exists(path)
        syms = subprocess.run(["nm", "-C", path], stdout=subprocess.PIPE, text=True).stdout
        # the reference's symbols are DEFINED by the adaptors, the C-ABI is imported from libsphb.so
        assert " T DensityCalculator::operator()(Particle&)" in syms
        assert " T setup_ghost_particles(" in syms
        assert " T SPHSimulation::step_forward()" in syms
        assert " U sphb_step" in syms and " U sphb_density" in syms


def test_no_cpu_fallback_in_the_dropin_program():
    """Without a CUDA device the program must fail loudly, not fall back to host arithmetic."""
    exe = os.path.join(ADAPTORS, "_build", "sph_dropin")
    if not os.path.exists(exe):
        pytest.skip("drop-in program not built")
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a CUDA device is present")
    except Exception:
        pass
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        with open(os.path.join(d, "config.txt"), "w") as f:
            f.write("n_part 101\nmass 0.04\npressure_calc 1\nlimit 2\nv_0 1\nh_factor 2\nt_i 0.005\n")
        res = subprocess.run([exe, "config.txt"], cwd=d, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert res.returncode != 0
    assert "no CPU fallback" in res.stdout or "CUDA" in res.stdout or "libsphb" in res.stdout
