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
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "basictypes.hpp"   // the reference's (sph/basictypes.hpp)
#include "define.hpp"       // the reference's (sph/define.hpp)
#include "sphb.h"

static_assert(sizeof(Particle) == sizeof(sphb_particle), "Particle layout (basictypes.hpp:49-116)");
static_assert(sizeof(Config) == sizeof(sphb_config), "Config layout (basictypes.hpp:21-31)");

namespace sphb_adaptor {

inline sphb_options default_options() {
    sphb_options o;
    sphb_default_options(&o);
#ifdef USE_VARIABLE_H
    o.variable_h = 1;       // define.hpp:27
#else
    o.variable_h = 0;
#endif
    o.fixed_h = CONSTANT_H; // define.hpp:29
    return o;
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
        return arr[i].h == image_[(size_t)i].h;
    }
    bool same_force_inputs(const Particle *arr, int i, bool own_pressure) const {
        for (int k = 0; k < n_; k++) {
            const sphb_particle &s = image_[(size_t)k];
            const Particle &a = arr[k];
            if (a.pos != s.pos || a.vel != s.vel || a.h != s.h || a.u != s.u || a.density != s.density ||
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
