// sphb_bridge.hpp — shared plumbing of the C++ adaptors: one libsphb handle behind the
// reference's per-particle functor protocol.
//
// The adaptors compile against the reference's OWN headers (-I<reference>/sph; nothing of
// the reference is copied here) and define the reference's symbols on top of the C-ABI of
// include/sphb.h, so the reference's main.cpp / setup.cpp link against them unchanged.
//
// A functor call (calculators.hpp:26,53,67,125) works on ONE particle of a host array the
// caller may have edited since the previous call.  The device pass is batch-shaped, so the
// bridge keeps the image of the array it last uploaded together with the batch results,
// and answers a call from those results while the INPUTS of that call are unchanged:
//   dc(p)      depends on pos, mass of everybody and on p's own h        
//   ac(p)/ec(p) depend on pos, vel, h, u, density, mass, type of everybody and p's own pressure
// Anything else (acc, du_dt, other particles' pressure) is output only.  When an input
// differs the array is uploaded again and the batch re-run, which is what makes the
// adaptors exact for ANY call pattern, including the reference's in-place ordered sweep
// (every kick invalidates the image: O(N) per call there, like the reference's own loops).
#pragma once
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <stdexcept>
#include <string>
#include <vector>

#include "basictypes.hpp"   // the reference's (sph/basictypes.hpp)
#include "define.hpp"       // the reference's (sph/define.hpp)
#include "sphb.h"

static_assert(sizeof(Particle) == sizeof(sphb_particle), "Particle layout (basictypes.hpp:49-116)");
static_assert(sizeof(Config) == sizeof(sphb_config), "Config layout (basictypes.hpp:21-31)");

namespace sphb_adaptor {

// Optional run-time keys, read from the SAME config file the reference's main.cpp reads (argv[1], else
// ./config.txt: main.cpp:29-35), in the reference's own "KEY VALUE" / "#" comment syntax (setup.cpp:44-89).
// The reference's ConfigReader only looks up the seven keys it knows (setup.cpp:27-33), so a config file carrying
// these extra keys still works with the reference, and a shipped config without them works here:
//   variable_h 0|1        per-particle root-find (1) or one fixed smoothing length (0); replaces the compile-time
//                         switch USE_VARIABLE_H (define.hpp:27) — one binary serves both
//   fixed_h <double>      that smoothing length (default CONSTANT_H, define.hpp:29)
//   ordering reference|jacobi     sweep order of the second half of the step (sphb_options.order)
//   sweep_rel_tol <double>        sphb_options.sweep_rel_tol
//   sweep_mode auto|worklist|serial   sphb_options.sweep_mode
//   dump_every <k>        file_write writes dumps/<step>.txt for every k-th step only (and the last one); the state is
//                         copied back from the device only for those steps
//   dump_stride <s>       ... and only every s-th alive particle of it (plus all ghosts)
// SPHB_CONFIG in the environment names the file when the program's argv is not the reference's.
struct RuntimeKeys {
    int variable_h = -1;          // -1: not given
    double fixed_h = -1.0;
    int order = -1;
    double sweep_rel_tol = -1.0;
    int sweep_mode = -1;
    int dump_every = 1, dump_stride = 1;
};

inline std::string config_path() {
    if (const char *e = std::getenv("SPHB_CONFIG")) return e;
    std::ifstream cmd("/proc/self/cmdline", std::ios::binary);   // argv, NUL-separated
    std::string arg0, arg1;
    if (cmd && std::getline(cmd, arg0, '\0') && std::getline(cmd, arg1, '\0') && !arg1.empty()) return arg1;
    return "./config.txt";
}

inline int key_choice(const std::string &v, const char *const *names, int n, const char *key) {
    for (int k = 0; k < n; k++)
        if (v == names[k] || v == std::to_string(k)) return k;
    std::cerr << "[ERROR] Failed to parse value '" << v << "' for property '" << key << "'." << std::endl;
    std::exit(1);   // what the reference's reader does with a value it cannot parse (setup.cpp:104-135)
}

inline const RuntimeKeys &runtime_keys() {
    static RuntimeKeys keys;
    static bool loaded = false;
    if (loaded) return keys;
    loaded = true;
    std::ifstream in(config_path());
    std::string l;
    while (in && std::getline(in, l)) {
        if (l.empty() || l[0] == '#') continue;
        const size_t space = l.find(' ');
        if (space == std::string::npos || space == 0 || space + 1 >= l.size()) continue;   // the reference's reader reports these
        const std::string name = l.substr(0, space);
        std::string value = l.substr(space + 1);
        while (!value.empty() && (value.back() == ' ' || value.back() == '\r' || value.back() == '\t')) value.pop_back();
        try {
            static const char *const orders[] = {"reference", "jacobi"};
            static const char *const modes[] = {"auto", "worklist", "serial"};
            if (name == "variable_h") keys.variable_h = std::stoi(value) != 0;
            else if (name == "fixed_h") keys.fixed_h = std::stod(value);
            else if (name == "ordering") keys.order = key_choice(value, orders, 2, "ordering");
            else if (name == "sweep_rel_tol") keys.sweep_rel_tol = std::stod(value);
            else if (name == "sweep_mode") keys.sweep_mode = key_choice(value, modes, 3, "sweep_mode");
            else if (name == "dump_every") keys.dump_every = std::max(1, std::stoi(value));
            else if (name == "dump_stride") keys.dump_stride = std::max(1, std::stoi(value));
        } catch (const std::exception &) {
            std::cerr << "[ERROR] Failed to parse value '" << value << "' for property '" << name << "'." << std::endl;
            std::exit(1);
        }
    }
    return keys;
}

inline sphb_options default_options() {
    sphb_options o;
    sphb_default_options(&o);
#ifdef USE_VARIABLE_H
    o.variable_h = 1;       // define.hpp:27
#else
    o.variable_h = 0;
#endif
    o.fixed_h = CONSTANT_H; // define.hpp:29
    const RuntimeKeys &k = runtime_keys();
    if (k.variable_h >= 0) o.variable_h = k.variable_h;
    if (k.fixed_h > 0) o.fixed_h = k.fixed_h;
    if (k.order >= 0) o.order = k.order;
    if (k.sweep_rel_tol >= 0) o.sweep_rel_tol = k.sweep_rel_tol;
    if (k.sweep_mode >= 0) o.sweep_mode = k.sweep_mode;
    return o;
}

// One fixed smoothing length for every particle?  (The reference's fixed-h build gives every particle CONSTANT_H in
// init_particles, setup.cpp:181-184.  With the run-time key the reference's unchanged init_particles has stored
// h_factor * spacing instead, so the adaptors put the fixed value in whenever they take the array to the device.)
inline bool fixed_mode() { return default_options().variable_h == 0; }
inline void impose_fixed_h(sphb_particle *rec, int n) {
    if (!fixed_mode()) return;
    const double h = default_options().fixed_h;
    for (int k = 0; k < n; k++) rec[k].h = h;
}

[[noreturn]] inline void fail(sphb_handle *h, int rc) {
    const std::string msg = sphb_last_error(h);
    // the reference throws these two as POINTERS (calculators.cpp:41,184)
    if (rc == SPHB_ERR_NEGATIVE_H || rc == SPHB_ERR_LOW_DENSITY) throw new std::logic_error(msg);
    if (rc == SPHB_ERR_GHOST_ZERO_H) throw std::logic_error(msg);   // ghost_particles.cpp:40
    throw std::runtime_error("libsphb: " + msg);
}

class Bridge {
public:
    static Bridge &get() { static Bridge b; return b; }

    // handle sized for at least n_alive particles
    sphb_handle *handle(const Config &c) {
        const int na = c.n_part - c.n_ghost;
        if (h_ && na <= cap_ && same_physics(c)) return h_;
        if (h_) sphb_destroy(h_);
        h_ = nullptr;
        sphb_options o = default_options();
        o.capacity = na + na / 8 + 64;
        sphb_config sc;
        std::memcpy(&sc, &c, sizeof sc);
        int rc = sphb_create(&sc, &o, &h_);
        if (rc) fail(nullptr, rc);
        cap_ = o.capacity;
        pad_ = o.pad_capacity;
        cfg_ = c;
        image_valid_ = false;
        return h_;
    }
    int max_records() const { return cap_ + 2 * pad_; }

    // --- DensityCalculator::operator()
    void density(const Config &c, Particle *arr, Particle &p) {
        const long i = index_of(c, arr, p);
        if (!(image_valid_ && dens_valid_ && same_layout(c) && same_density_inputs(arr, (int)i))) {
            upload(c, arr);
            int rc = sphb_density(h_, 1);
            if (rc) { image_valid_ = false; fail(h_, rc); }
            download(dens_);
            dens_valid_ = true;
        }
        p.h = dens_[(size_t)i].h;
        p.density = dens_[(size_t)i].density;
    }

    // --- AccelerationCalculator / EnergyCalculator ::operator()
    const sphb_particle &forces(const Config &c, Particle *arr, Particle &p, bool own_pressure_is_input) {
        const long i = index_of(c, arr, p);
        if (!(image_valid_ && forc_valid_ && same_layout(c) && same_force_inputs(arr, (int)i, own_pressure_is_input))) {
            upload(c, arr);
            int rc = sphb_forces(h_);
            if (rc) { image_valid_ = false; fail(h_, rc); }
            download(forc_);
            forc_valid_ = true;
        }
        return forc_[(size_t)i];
    }

    void invalidate() { image_valid_ = false; }

private:
    Bridge() = default;
    ~Bridge() { if (h_) sphb_destroy(h_); }

    static long index_of(const Config &c, Particle *arr, Particle &p) {
        const long i = &p - arr;
        if (i < 0 || i >= c.n_part)
            throw std::logic_error("libsphb adaptor: the particle is not an element of the bound particle array");
        return i;
    }
    bool same_physics(const Config &c) const {
        return c.mass == cfg_.mass && c.pressure_calc == cfg_.pressure_calc && c.limit == cfg_.limit &&
               c.v_0 == cfg_.v_0 && c.h_factor == cfg_.h_factor && c.t_i == cfg_.t_i;
    }
    bool same_layout(const Config &c) const { return c.n_part == n_ && c.n_ghost == g_ && same_physics(c); }
    bool same_density_inputs(const Particle *arr, int i) const {
        for (int k = 0; k < n_; k++)
            if (arr[k].pos != image_[(size_t)k].pos || arr[k].mass != image_[(size_t)k].mass) return false;
        return fixed_mode() || arr[i].h == image_[(size_t)i].h;
    }
    bool same_force_inputs(const Particle *arr, int i, bool own_pressure) const {
        for (int k = 0; k < n_; k++) {
            const sphb_particle &s = image_[(size_t)k];
            const Particle &a = arr[k];
            if (a.pos != s.pos || a.vel != s.vel || (a.h != s.h && !fixed_mode()) || a.u != s.u || a.density != s.density ||
                a.mass != s.mass || (int)a.type != s.type)
                return false;
        }
        return !own_pressure || arr[i].pressure == image_[(size_t)i].pressure;
    }
    void upload(const Config &c, const Particle *arr) {
        handle(c);
        n_ = c.n_part;
        g_ = c.n_ghost;
        image_.resize((size_t)n_);
        std::memcpy(image_.data(), arr, sizeof(sphb_particle) * (size_t)n_);
        impose_fixed_h(image_.data(), n_);
        int rc = sphb_upload(h_, image_.data(), n_, g_);
        if (rc) { image_valid_ = false; fail(h_, rc); }
        image_valid_ = true;
        dens_valid_ = forc_valid_ = false;
    }
    void download(std::vector<sphb_particle> &out) {
        out.resize((size_t)max_records());
        int n = 0, g = 0;
        int rc = sphb_download(h_, out.data(), (int)out.size(), &n, &g, 1);
        if (rc) fail(h_, rc);
        out.resize((size_t)n);
    }

    sphb_handle *h_ = nullptr;
    Config cfg_{};
    int cap_ = 0, pad_ = 0, n_ = 0, g_ = 0;
    bool image_valid_ = false, dens_valid_ = false, forc_valid_ = false;
    std::vector<sphb_particle> image_, dens_, forc_;
};

}  // namespace sphb_adaptor
