// sph_simulation_sphb.cpp — SPHSimulation of the reference's sph/sph_simulation.hpp with
// step_forward() running on the GPU (replaces sph/sph_simulation.cpp).  The state lives in
// HBM between steps; it is copied back after each step only because file_write() dumps it.
#include <cstdio>
#include <filesystem>
#include <map>

#include "sph_simulation.hpp"   // the reference's
#include "sphb_bridge.hpp"

namespace {
struct DeviceSide {
    sphb_handle *h = nullptr;
    int host_slots = 0;   // records the host array can hold
};
std::map<const SPHSimulation *, DeviceSide> g_device;   // the reference's class has no spare member
}  // namespace

void SPHSimulation::start(double end_time) {
    std::cout << "[INFO] Simulation time: " << current_time << " / " << end_time << std::endl;
    std::error_code dir_ec;
    if (!std::filesystem::exists("dumps") && !std::filesystem::create_directory("dumps", dir_ec)) {
        std::cerr << "[ERROR] Failed to make directory ./dumps/ to store dump files: " << dir_ec.message() << std::endl;
        exit(1);
    }
    file_write();
#ifndef SETUP_ONLY
    while (current_time < (end_time - CALC_EPSILON)) {   // sph_simulation.cpp:44
        current_time += timestep;
        std::cout << "[INFO] Simulation time: " << current_time << " / " << end_time << std::endl;
        step_forward();
        file_write();
    }
#endif
    auto it = g_device.find(this);
    if (it != g_device.end()) { sphb_destroy(it->second.h); g_device.erase(it); }
}

// sph_simulation.cpp:55-108, on the device
void SPHSimulation::step_forward() {
    using namespace sphb_adaptor;
    DeviceSide &d = g_device[this];
    if (!d.h) {
        sphb_options o = default_options();
        sphb_config sc;
        std::memcpy(&sc, &config, sizeof sc);
        int rc = sphb_create(&sc, &o, &d.h);
        if (rc) fail(nullptr, rc);
        rc = sphb_upload(d.h, reinterpret_cast<const sphb_particle *>(p_arr.get()), config.n_part, config.n_ghost);
        if (rc) fail(d.h, rc);
        d.host_slots = config.n_part;
    }
    int rc = sphb_step(d.h, 1);
    if (rc) fail(d.h, rc);
    sphb_config now;
    sphb_get_config(d.h, &now);
    if (now.n_part > d.host_slots) {   // more ghosts than before: grow the host array
        ParticleArrayPtr grown(new Particle[now.n_part]);
        p_arr.swap(grown);
        d.host_slots = now.n_part;
    }
    int n_part = 0, n_ghost = 0;
    rc = sphb_download(d.h, reinterpret_cast<sphb_particle *>(p_arr.get()), d.host_slots, &n_part, &n_ghost, 0);
    if (rc) fail(d.h, rc);
    config.n_part = n_part;
    config.n_ghost = n_ghost;
    dc.update(config, p_arr);
    ac.update(config, p_arr);
    ec.update(config, p_arr);
}

// sph_simulation.cpp:110-139: same file names, header lines and column format, so that the
// reference's plot.py reads the dumps unchanged
void SPHSimulation::file_write() {
    outstream.open("./dumps/" + std::to_string(dump_counter) + ".txt");
    outstream << "# This file was dumped at t = " << current_time << std::endl
              << "# Column definitions:" << std::endl
              << "# Particle ID / Type / Smoothing length / Density / Pressure / Acceleration / Velocity / Position / Thermal energy" << std::endl
              << "# Aligned definition 'tags' for easier reading:" << std::endl
              << "# ID    TYPE     H          DENSITY  PRESS    ACCEL     VEL       POS       U" << std::endl;
    char line[256];
    for (int i = 0; i < config.n_part; i++) {
        const Particle &p = p_arr[i];
        std::snprintf(line, sizeof line, "%4d    %s    %3.5f    %3.3f    %3.3f    %+3.3f    %+3.3f    %+3.3f    %3.3f\n",
                      p.id, ParticleTypeNames[p.type], p.h, p.density, p.pressure, p.acc, p.vel, p.pos, p.u);
        outstream << line;
    }
    outstream.close();
    dump_counter++;
}
