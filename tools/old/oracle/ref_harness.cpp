/*
 * oracle/ref_harness.cpp — TEST INFRASTRUCTURE, not product code.
 *
 * A C-callable driver around the reference's OWN sources, which are compiled
 * unmodified from /root/reference/sph/*.cpp by oracle/Makefile into
 * oracle/_ref/ (never copied into this repository).  It exposes, for ctypes:
 *   - the full reference run: init_particles (reference sph/setup.cpp:143-228)
 *     followed by SPHSimulation::step_forward (sph/sph_simulation.cpp:55-108),
 *     with the complete double-precision state readable after every step
 *     (the reference's own dumps print 3-5 decimals, sph_simulation.cpp:131-133);
 *   - function-level calls of the reference functors on a caller-supplied
 *     state: DensityCalculator / AccelerationCalculator / EnergyCalculator
 *     (sph/calculators.cpp:33-46,63-129,192-213), calc_omega / rootfind_h
 *     (sph/smoothing_length.cpp:64-77,191-259), setup_ghost_particles
 *     (sph/ghost_particles.cpp:19-115);
 *   - the Newton iteration log kept by the GSL shim (oracle/gsl_shim).
 *
 * step_forward, config and p_arr are private members of SPHSimulation; the
 * reference header is included under "#define private public" (after every
 * standard header it needs has already been included) so the reference
 * sources stay untouched.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load the library this file builds.
 */
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <fstream>
#include <iostream>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include <gsl/gsl_roots.h>

#define private public
#define protected public
#include "sph_simulation.hpp"
#undef protected
#undef private
#include "setup.hpp"
#include "ghost_particles.hpp"
#include "smoothing_length.hpp"
#include "kernel.hpp"

#ifdef SPHB_ORACLE_FIXED_H
extern "C" double sphb_oracle_constant_h;
double sphb_oracle_constant_h = 0.2;
#endif

namespace {

/* Slack kept behind every particle array: ghost_particles.cpp:17 keeps a
 * file-static high-water mark that survives across simulations in one
 * process, so a later, smaller array would otherwise be overrun. */
const int kGhostSlack = 1 << 14;

struct NullBuf : std::streambuf {
    int overflow(int c) override { return c; }
};
NullBuf g_nullbuf;
std::streambuf *g_saved_cout = nullptr;

void quiet(bool on) {
    if (on && !g_saved_cout) g_saved_cout = std::cout.rdbuf(&g_nullbuf);
    if (!on && g_saved_cout) { std::cout.rdbuf(g_saved_cout); g_saved_cout = nullptr; }
}

ParticleArrayPtr make_array(int n) {
    /* zero-filled storage so that members the reference never initialises
     * (u in isothermal runs, pressure before the first force pass) are
     * deterministic zeros instead of heap garbage */
    size_t bytes = sizeof(Particle) * (size_t)(n + kGhostSlack);
    void *mem = std::calloc(1, bytes);
    if (!mem) throw std::bad_alloc();
    Particle *arr = new (mem) Particle[n + kGhostSlack];
    return ParticleArrayPtr(arr, [](Particle *p) { std::free((void *)p); });
}

struct Handle {
    Config config;
    ParticleArrayPtr p_arr;
    std::unique_ptr<SPHSimulation> sim;   /* only for full runs */
    std::string error;
    Config &cfg() { return sim ? sim->config : config; }
    ParticleArrayPtr &arr() { return sim ? sim->p_arr : p_arr; }
};

Config make_config(int n_part, double mass, int pressure_calc, double limit, double v_0,
                   double h_factor, double t_i) {
    Config c;
    std::memset((void *)&c, 0, sizeof c);
    c.n_part = n_part;
    c.mass = mass;
    c.pressure_calc = (PressureCalc)pressure_calc;
    c.limit = limit;
    c.v_0 = v_0;
    c.h_factor = h_factor;
    c.t_i = t_i;
    c.n_ghost = 0;
    return c;
}

template <class F>
int guarded(Handle *h, F &&f) {
    try {
        f();
        return 0;
    } catch (std::logic_error *e) {           /* the reference throws pointers */
        h->error = e->what();
        return 2;
    } catch (const std::exception &e) {
        h->error = e.what();
        return 1;
    }
}

}  // namespace

extern "C" {

int ref_sizeof_particle() { return (int)sizeof(Particle); }
int ref_sizeof_config() { return (int)sizeof(Config); }

int ref_is_variable_h() {
#ifdef USE_VARIABLE_H
    return 1;
#else
    return 0;
#endif
}

void ref_set_constant_h(double h) {
#ifdef SPHB_ORACLE_FIXED_H
    sphb_oracle_constant_h = h;
#else
    (void)h;
#endif
}

void ref_set_quiet(int on) { quiet(on != 0); }

/* Full reference start-up: lattice, adiabatic pre-pass, ghosts, t = 0 state. */
void *ref_create(int n_part, double mass, int pressure_calc, double limit, double v_0,
                 double h_factor, double t_i) {
    quiet(std::getenv("SPHB_ORACLE_VERBOSE") == nullptr);
    Handle *h = new Handle();
    h->config = make_config(n_part, mass, pressure_calc, limit, v_0, h_factor, t_i);
    h->p_arr = make_array(n_part);
    int rc = guarded(h, [&] {
        init_particles(h->config, h->p_arr);
        h->sim.reset(new SPHSimulation(h->config, h->p_arr));
    });
    (void)rc;
    return h;
}

/* A handle over a caller-supplied state, no init_particles.  fields is
 * [n_total][9] = mass,pos,vel,acc,h,du_dt,u,density,pressure; type[n_total]. */
void *ref_create_raw(int n_alive, int n_ghost, double mass, int pressure_calc, double limit,
                     double v_0, double h_factor, double t_i, const double *fields,
                     const int *type) {
    quiet(std::getenv("SPHB_ORACLE_VERBOSE") == nullptr);
    Handle *h = new Handle();
    int n_total = n_alive + n_ghost;
    h->config = make_config(n_total, mass, pressure_calc, limit, v_0, h_factor, t_i);
    h->config.n_ghost = n_ghost;
    h->p_arr = make_array(n_total);
    for (int i = 0; i < n_total; i++) {
        Particle &p = h->p_arr[i];
        const double *f = fields + 9 * (size_t)i;
        p.mass = f[0]; p.pos = f[1]; p.vel = f[2]; p.acc = f[3]; p.h = f[4];
        p.du_dt = f[5]; p.u = f[6]; p.density = f[7]; p.pressure = f[8];
        p.type = (ParticleType)type[i];
    }
    return h;
}

void ref_destroy(void *vh) { delete (Handle *)vh; }

const char *ref_last_error(void *vh) { return ((Handle *)vh)->error.c_str(); }

int ref_n_part(void *vh) { return ((Handle *)vh)->cfg().n_part; }
int ref_n_ghost(void *vh) { return ((Handle *)vh)->cfg().n_ghost; }

/* SPHSimulation::step_forward, nsteps times (sph_simulation.cpp:55-108). */
int ref_step(void *vh, int nsteps) {
    Handle *h = (Handle *)vh;
    if (!h->sim) {
        /* promote a raw handle to a simulation on its current state */
        h->sim.reset(new SPHSimulation(h->config, h->p_arr));
    }
    return guarded(h, [&] {
        for (int s = 0; s < nsteps; s++) {
            h->sim->current_time += h->sim->timestep;
            h->sim->step_forward();
        }
    });
}

void ref_get_state(void *vh, double *fields, int *type, int *id) {
    Handle *h = (Handle *)vh;
    int n = h->cfg().n_part;
    ParticleArrayPtr &a = h->arr();
    for (int i = 0; i < n; i++) {
        const Particle &p = a[i];
        double *f = fields + 9 * (size_t)i;
        f[0] = p.mass; f[1] = p.pos; f[2] = p.vel; f[3] = p.acc; f[4] = p.h;
        f[5] = p.du_dt; f[6] = p.u; f[7] = p.density; f[8] = p.pressure;
        if (type) type[i] = (int)p.type;
        if (id) id[i] = p.id;
    }
}

/* ---- function-level calls on the handle's current state ---- */

int ref_op_ghosts(void *vh) {
    Handle *h = (Handle *)vh;
    return guarded(h, [&] {
        setup_ghost_particles(h->arr(), h->cfg());
        if (h->sim) {
            h->sim->dc.update(h->cfg(), h->arr());
            h->sim->ac.update(h->cfg(), h->arr());
            h->sim->ec.update(h->cfg(), h->arr());
        }
    });
}

/* dc(p) for i in [i0, i1), in order, in place. */
int ref_op_density(void *vh, int i0, int i1) {
    Handle *h = (Handle *)vh;
    return guarded(h, [&] {
        DensityCalculator dc(h->cfg(), h->arr());
        for (int i = i0; i < i1; i++) dc(h->arr()[i]);
    });
}

/* ac(p); ec(p) for i in [i0, i1), in order, in place, no kick
 * (the t = 0 pass of setup.cpp:223-226). */
int ref_op_forces(void *vh, int i0, int i1) {
    Handle *h = (Handle *)vh;
    return guarded(h, [&] {
        AccelerationCalculator ac(h->cfg(), h->arr());
        EnergyCalculator ec(h->cfg(), h->arr());
        for (int i = i0; i < i1; i++) {
            ac(h->arr()[i]);
            ec(h->arr()[i]);
        }
    });
}

/* first loop of step_forward only (sph_simulation.cpp:59-73) */
int ref_op_kick_drift(void *vh) {
    Handle *h = (Handle *)vh;
    Config &c = h->cfg();
    double dt = c.t_i;
    for (int i = 0; i < c.n_part; i++) {
        Particle &p = h->arr()[i];
        if (p.type == Ghost) continue;
        p.vel += p.acc * (dt / 2);
        p.u += p.du_dt * (dt / 2);
        p.pos += p.vel * (dt);
    }
    return 0;
}

/* calc_omega for every particle with its stored h, density */
int ref_op_omega(void *vh, double *omega_out) {
    Handle *h = (Handle *)vh;
    return guarded(h, [&] {
        int n = h->cfg().n_part;
        for (int i = 0; i < n; i++) omega_out[i] = calc_omega(h->arr()[i], h->arr(), h->cfg());
    });
}

double ref_kernel(double q) { return kernel(q); }
double ref_dkernel_dq(double q) { return dkernel_dq(q); }

/* Newton log: one entry per rootfind_h call since the last read. */
int ref_newton_log(int *iters_out, int cap) {
    int n = (int)sphb_shim_log.newton_iters.size();
    int m = std::min(n, cap);
    for (int i = 0; i < m; i++) iters_out[i] = sphb_shim_log.newton_iters[i];
    sphb_shim_log.newton_iters.clear();
    return n;
}
int ref_bisect_log_count() {
    int n = (int)sphb_shim_log.bisect_iters.size();
    sphb_shim_log.bisect_iters.clear();
    return n;
}

}  // extern "C"
