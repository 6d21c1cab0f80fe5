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


KEYS_PROBE = r"""
#include <cstdio>
#include "sphb_bridge.hpp"
int main() {
    const sphb_options o = sphb_adaptor::default_options();
    const sphb_adaptor::RuntimeKeys &k = sphb_adaptor::runtime_keys();
    std::printf("%d %.3f %d %.1e %d %d %d %d\n", o.variable_h, o.fixed_h, o.order, o.sweep_rel_tol, o.sweep_mode,
                k.dump_every, k.dump_stride, (int)sphb_adaptor::fixed_mode());
    return 0;
}
"""


@pytest.mark.skipif(not os.path.isdir("/root/reference/sph"), reason="reference sources not present")
def test_adaptors_read_optional_keys_from_the_reference_config_file(tmp_path):
    """The adaptors' run-time keys (sphb_bridge.hpp) in the reference's own config syntax: a shipped config leaves the
    compile-time defaults, an extended one overrides them, and the file is found the way main.cpp finds it (argv[1])."""
    import sph_b200.build as b
    b.build()
    src = tmp_path / "probe.cpp"
    src.write_text(KEYS_PROBE)
    exe = tmp_path / "probe"
    cmd = ["g++", "-std=c++17", "-w", "-I/root/reference/sph", "-I" + os.path.join(ROOT, "include"), "-I" + ADAPTORS,
           "-o", str(exe), str(src), "-L" + os.path.join(ROOT, "sph_b200"), "-lsphb",
           "-Wl,-rpath," + os.path.join(ROOT, "sph_b200")]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert res.returncode == 0, res.stdout[-2000:]
    shipped = "# comment\nn_part 101\nmass 0.04\npressure_calc 1\nlimit 2\nv_0 1\nh_factor 2\nt_i 0.005\n"
    (tmp_path / "a.txt").write_text(shipped)
    (tmp_path / "b.txt").write_text(shipped + "variable_h 0\nfixed_h 0.125\nordering jacobi\nsweep_rel_tol 1e-12\n"
                                    "sweep_mode serial\ndump_every 50\ndump_stride 10\n# variable_h 1\nunknown_key 7\n")
    env = {k: v for k, v in os.environ.items() if k != "SPHB_CONFIG"}
    out = subprocess.run([str(exe), "a.txt"], cwd=tmp_path, env=env, stdout=subprocess.PIPE, text=True).stdout.split()
    assert out == ["1", "0.200", "0", "0.0e+00", "0", "1", "1", "0"], out
    out = subprocess.run([str(exe), "b.txt"], cwd=tmp_path, env=env, stdout=subprocess.PIPE, text=True).stdout.split()
    assert out == ["0", "0.125", "1", "1.0e-12", "2", "50", "10", "1"], out
    # SPHB_CONFIG names the file when the program's argv is not the reference's
    out = subprocess.run([str(exe)], cwd=tmp_path, env=dict(env, SPHB_CONFIG=str(tmp_path / "b.txt")),
                         stdout=subprocess.PIPE, text=True).stdout.split()
    assert out[0] == "0" and out[5] == "50"
    # a value the reader cannot parse ends the program like the reference's reader does (exit 1, setup.cpp:104-135)
    (tmp_path / "c.txt").write_text(shipped + "ordering sideways\n")
    res = subprocess.run([str(exe), "c.txt"], cwd=tmp_path, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert res.returncode == 1 and "ordering" in res.stdout
