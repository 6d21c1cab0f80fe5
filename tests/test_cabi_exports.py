"""CPU tests of the boundary: the C-ABI library loads, exports every symbol that
include/sphb.h declares, mirrors the reference's struct layouts, and FAILS LOUDLY without a
CUDA device (there is no CPU fallback to fall into)."""
import ctypes as C

import numpy as np
import pytest

import sph_b200
from sph_b200 import api


def test_every_declared_symbol_is_exported():
    declared = api.declared_symbols()
    assert len(declared) >= 20
    assert "sphb_step" in declared and "sphb_density" in declared
    assert set(api.exported_symbols()) == set(declared)


def test_struct_layouts_match_the_reference():
    # sizeof(Particle) = 88, sizeof(Config) = 64 (probed on the reference)
    assert api.PARTICLE_DTYPE.itemsize == 88
    assert C.sizeof(api.SphbConfig) == 64
    assert api.SphbConfig.mass.offset == 8 and api.SphbConfig.n_ghost.offset == 56
    off = {n: api.PARTICLE_DTYPE.fields[n][1] for n in api.PARTICLE_DTYPE.names}
    assert off == {"id": 0, "mass": 8, "pos": 16, "vel": 24, "acc": 32, "h": 40, "du_dt": 48,
                   "u": 56, "density": 64, "pressure": 72, "type": 80}


def test_build_arch():
    assert api.load_library().sphb_build_arch() == b"sm_100a"


def test_no_cpu_fallback(cuda_available):
    if cuda_available:
        pytest.skip("a CUDA device is present")
    with pytest.raises(sph_b200.SphbError) as e:
        sph_b200.Simulation(sph_b200.Config(101, 0.04, 1))
    assert e.value.code in (2, 12)
    assert "no CPU fallback" in str(e.value) or "CUDA" in str(e.value)


def test_invalid_arguments_are_rejected_before_touching_the_device():
    L = api.load_library()
    h = C.c_void_p()
    cfg = api.SphbConfig(0, 0.04, 1, 2.0, 1.0, 2.0, 0.005, 0)
    assert L.sphb_create(C.byref(cfg), None, C.byref(h)) == 1
    cfg = api.SphbConfig(101, 0.04, 7, 2.0, 1.0, 2.0, 0.005, 0)
    assert L.sphb_create(C.byref(cfg), None, C.byref(h)) == 1
    assert b"Unknown pressure calculation mode" in L.sphb_last_error(None)


def test_ctypes_structs_match_the_c_header(tmp_path):
    """include/sphb.h compiles as plain C99 and its struct sizes are the ones the ctypes binding assumes."""
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include "sphb.h"\nint main(void){printf("%zu %zu %zu %zu\\n", sizeof(sphb_config), '
                   'sizeof(sphb_particle), sizeof(sphb_options), sizeof(sphb_counters));return 0;}\n')
    exe = tmp_path / "sizes"
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror",
                           "-I" + os.path.join(root, "include"), str(src), "-o", str(exe)])
    sizes = [int(v) for v in subprocess.check_output([str(exe)]).decode().split()]
    assert sizes == [C.sizeof(api.SphbConfig), api.PARTICLE_DTYPE.itemsize, C.sizeof(api.SphbOptions),
                     C.sizeof(api.SphbCounters)]
