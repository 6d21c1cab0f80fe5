// smoothing_length_sphb.cpp — the free functions of the reference's sph/smoothing_length.hpp
// on top of libsphb (replaces sph/smoothing_length.cpp; no GSL needed any more).
#include <vector>

#include "smoothing_length.hpp"   // the reference's
#include "sphb_bridge.hpp"

namespace {
// run one batch pass over a scratch copy of the array and return the record of particle i
sphb_particle batch_density_record(const Particle *arr, const Config &c, long i, double h_override, bool root) {
    std::vector<sphb_particle> img((size_t)c.n_part);
    std::memcpy(img.data(), arr, sizeof(sphb_particle) * (size_t)c.n_part);
    if (h_override > 0) img[(size_t)i].h = h_override;
    sphb_handle *h = nullptr;
    sphb_options o = sphb_adaptor::default_options();
    if (!root) o.variable_h = 0;          // density sum at the stored h, no root-find
    sphb_config sc;
    std::memcpy(&sc, &c, sizeof sc);
    int rc = sphb_create(&sc, &o, &h);
    if (rc) sphb_adaptor::fail(nullptr, rc);
    rc = sphb_upload(h, img.data(), c.n_part, c.n_ghost);
    if (!rc) rc = sphb_density(h, 1);
    int n = 0, g = 0;
    std::vector<sphb_particle> out((size_t)c.n_part + 2 * (size_t)o.pad_capacity);
    if (!rc) rc = sphb_download(h, out.data(), (int)out.size(), &n, &g, 1);
    if (rc) { std::string m = sphb_last_error(h); sphb_destroy(h); throw std::runtime_error("libsphb: " + m); }
    sphb_destroy(h);
    return out[(size_t)i];
}
long find(const Particle &p, const Particle *arr, int n) {
    for (int k = 0; k < n; k++) if (arr[k].id == p.id) return k;
    throw std::logic_error("libsphb adaptor: particle not found in the array (by id)");
}
}  // namespace

// smoothing_length.cpp:80-91
double calc_density(const Particle &p, const double h, const Particle *p_arr, const int n_part) {
    Config c{};
    c.n_part = n_part; c.n_ghost = 0; c.mass = p.mass; c.pressure_calc = Isothermal;
    c.limit = 1; c.v_0 = 0; c.h_factor = 1; c.t_i = 1;
    return batch_density_record(p_arr, c, find(p, p_arr, n_part), h, false).density;
}

// smoothing_length.cpp:191-259
double rootfind_h(const Particle &p, const ParticleArrayPtr p_arr, const Config c) {
    return batch_density_record(p_arr.get(), c, find(p, p_arr.get(), c.n_part), -1.0, true).h;
}

// smoothing_length.cpp:64-77: 1 + (h/rho) sum_j m_j dW/dh at the stored h, rho
double calc_omega(const Particle &p, ParticleArrayPtr p_arr, Config c) {
#ifndef USE_VARIABLE_H
    (void)p; (void)p_arr; (void)c;
    return 1;
#else
    if (sphb_adaptor::fixed_mode()) return 1;    // `variable_h 0` in the config file
    const long i = find(p, p_arr.get(), c.n_part);
    sphb_handle *h = nullptr;
    sphb_options o = sphb_adaptor::default_options();
    sphb_config sc;
    std::memcpy(&sc, &c, sizeof sc);
    int rc = sphb_create(&sc, &o, &h);
    if (rc) sphb_adaptor::fail(nullptr, rc);
    std::vector<double> om((size_t)c.n_part + 2 * (size_t)o.pad_capacity);
    rc = sphb_upload(h, reinterpret_cast<const sphb_particle *>(p_arr.get()), c.n_part, c.n_ghost);
    if (!rc) rc = sphb_omega(h, om.data(), (int)om.size());
    if (rc) { std::string m = sphb_last_error(h); sphb_destroy(h); throw std::runtime_error("libsphb: " + m); }
    sphb_destroy(h);
    return om[(size_t)i];
#endif
}
