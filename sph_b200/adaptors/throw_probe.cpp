// throw_probe.cpp — test program (tests/test_dropin_binary.py): do the adaptors rethrow device errors the way the
// reference throws them?  calculators.cpp:184 throws `new std::logic_error` (a POINTER) for a density below epsilon,
// ghost_particles.cpp:39 throws std::logic_error by value for a zero smoothing length.  A host program written against
// the reference catches exactly those types.  Exit code 0 = both arrived with the right type and message.
#include <cstdio>
#include <cstring>
#include <stdexcept>

#include "calculators.hpp"       // the reference's
#include "ghost_particles.hpp"   // the reference's

int main() {
    const int n = 50;
    Config c;
    std::memset((void *)&c, 0, sizeof c);
    c.n_part = n; c.mass = 1e-12; c.pressure_calc = Adiabatic; c.limit = 2; c.v_0 = 1; c.h_factor = 2; c.t_i = 0.005; c.n_ghost = 0;
    ParticleArrayPtr arr(new Particle[n]);
    for (int i = 0; i < n; i++) {
        Particle &p = arr[i];
        p.mass = c.mass; p.pos = -1.9 + 3.8 * i / (n - 1); p.vel = 0; p.acc = 0; p.h = 0.16; p.du_dt = 0; p.u = 1.5;
        p.density = 1; p.pressure = 1;
    }
    int got = 0;
    try {
        DensityCalculator dc(c, arr);
        AccelerationCalculator ac(c, arr);
        for (int i = 0; i < n; i++) dc(arr[i]);
        ac(arr[0]);
        std::printf("no exception for densities of 1e-11\n");
    } catch (std::logic_error *e) {
        std::printf("pointer: %s\n", e->what());
        if (std::strstr(e->what(), "density less than epsilon")) got |= 1;
    } catch (const std::exception &e) {
        std::printf("value (wrong): %s\n", e.what());
    }
    for (int i = 0; i < n; i++) { arr[i].mass = 0.08; arr[i].h = 0; }
    c.mass = 0.08;
    try {
        setup_ghost_particles(arr, c);
        std::printf("no exception for zero smoothing lengths\n");
    } catch (std::logic_error *e) {
        std::printf("pointer (wrong): %s\n", e->what());
    } catch (const std::logic_error &e) {
        std::printf("value: %s\n", e.what());
        if (std::strstr(e.what(), "zero smoothing length")) got |= 2;
    } catch (const std::exception &e) {
        std::printf("other (wrong): %s\n", e.what());
    }
    return got == 3 ? 0 : 1;
}
