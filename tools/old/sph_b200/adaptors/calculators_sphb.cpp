// calculators_sphb.cpp — defines the functors of the reference's sph/calculators.hpp on top
// of libsphb (replaces sph/calculators.cpp).  Compiled against the reference's own header.
#include <cmath>

#include "calculators.hpp"   // the reference's
#include "sphb_bridge.hpp"

using sphb_adaptor::Bridge;

// calculators.cpp:33-55 — p.h <- rootfind_h(p); p.density <- calc_density(p, p.h)
void DensityCalculator::operator()(Particle &p) {
    Bridge::get().density(config, p_arr.get(), p);
}

// calculators.cpp:63-129 — ghosts are skipped; p.pressure and p.acc are set
void AccelerationCalculator::operator()(Particle &p_i) {
    if (p_i.type == Ghost) return;
    if (config.pressure_calc != Isothermal && config.pressure_calc != Adiabatic)
        throw std::logic_error("Unknown pressure calculation mode!");
    const sphb_particle &r = Bridge::get().forces(config, p_arr.get(), p_i, /*own pressure feeds c_s*/ true);
    p_i.pressure = r.pressure;
    p_i.acc = r.acc;
}

// calculators.cpp:192-213 — p.du_dt from the pressure ac(p) stored just before
void EnergyCalculator::operator()(Particle &p) {
    const sphb_particle &r = Bridge::get().forces(config, p_arr.get(), p, false);
    p.du_dt = r.du_dt;
}

// calculators.cpp:139-146 — one record, host arithmetic
double AccelerationCalculator::sound_speed(Particle p) {
    if (config.pressure_calc == Isothermal) return 1;
    if (config.pressure_calc == Adiabatic) return std::sqrt(GAMMA * p.pressure / p.density);
    throw std::logic_error("Unknown pressure calculation mode!");
}
