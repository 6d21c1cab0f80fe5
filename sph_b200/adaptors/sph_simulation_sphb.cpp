// sph_simulation_sphb.cpp — SPHSimulation of the reference's sph/sph_simulation.hpp with
// step_forward() running on the GPU (replaces sph/sph_simulation.cpp).  The state lives in
// HBM between steps; it is copied back only for the steps file_write() dumps (every step by
// default, like the reference; `dump_every k` / `dump_stride s` in the config file thin that out).
#include <cstdio>
#include <filesystem>
#include <map>
#include <vector>

#include "sph_simulation.hpp"   // the reference's
#include "sphb_bridge.hpp"

namespace {
struct DeviceSide {
    sphb_handle *h = nullptr;
    int host_slots = 0;          // records the host array can hold
    bool host_current = true;    // p_arr holds the state of the latest step
    bool dumped_latest = false;  // ... and file_write has written it
    std::vector<sphb_particle> sample;   // dump_stride > 1: every s-th alive record + the ghosts of the latest dump step
    int sample_n = 0;
};
std::map<const SPHSimulation *, DeviceSide> g_device;   // the reference's class has no spare member
}  // namespace

namespace {
bool force_dump = false;

// the whole state back into the reference's host array (grown when there are more ghosts than before)
void sync_host_array(DeviceSide &d, Config &config, ParticleArrayPtr &p_arr, DensityCalculator &dc,
                     AccelerationCalculator &ac, EnergyCalculator &ec) {
    using namespace sphb_adaptor;
    sphb_config now;
    sphb_get_config(d.h, &now);
    if (now.n_part > d.host_slots) {
        ParticleArrayPtr grown(new Particle[now.n_part]);
        p_arr.swap(grown);
        d.host_slots = now.n_part;
    }
    int n_part = 0, n_ghost = 0;
    int rc = sphb_download(d.h, reinterpret_cast<sphb_particle *>(p_arr.get()), d.host_slots, &n_part, &n_ghost, 0);
    if (rc) fail(d.h, rc);
    config.n_part = n_part;
    config.n_ghost = n_ghost;
    dc.update(config, p_arr);
    ac.update(config, p_arr);
    ec.update(config, p_arr);
    d.host_current = true;
}
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
    if (it != g_device.end()) {
        // the caller finds the final state in p_arr, as after the reference's loop, and the last step is always dumped
        DeviceSide &d = it->second;
        if (d.h && !d.host_current) sync_host_array(d, config, p_arr, dc, ac, ec);
        if (d.h && !d.dumped_latest) { dump_counter--; d.sample_n = 0; force_dump = true; file_write(); force_dump = false; }
        sphb_destroy(d.h);
        g_device.erase(it);
    }
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
    d.host_current = false;
    d.dumped_latest = false;
    d.sample_n = 0;
    const RuntimeKeys &keys = runtime_keys();
    if (dump_counter % keys.dump_every != 0) return;      // file_write() will skip this step: the state stays on the device
    if (keys.dump_stride > 1) {                            // a thinned-out snapshot is all the dump needs
        sphb_config now;
        sphb_get_config(d.h, &now);
        const int na = now.n_part - now.n_ghost;
        d.sample.resize((size_t)(na / keys.dump_stride + 1 + 2 * default_options().pad_capacity));
        int na_out = 0, ng = 0;
        rc = sphb_download_strided(d.h, d.sample.data(), (int)d.sample.size(), keys.dump_stride, &na_out, &ng, 1);
        if (rc) fail(d.h, rc);
        d.sample_n = na_out + ng;
        return;
    }
    sync_host_array(d, config, p_arr, dc, ac, ec);
}

// sph_simulation.cpp:110-139: same file names, header lines and column format, so that the
// reference's plot.py reads the dumps unchanged
void SPHSimulation::file_write() {
    auto dit = g_device.find(this);
    if (!force_dump && dump_counter % sphb_adaptor::runtime_keys().dump_every != 0) {   // `dump_every k`: not this step
        dump_counter++;
        return;
    }
    if (dit != g_device.end()) dit->second.dumped_latest = true;
    const sphb_particle *sample = (dit != g_device.end() && dit->second.sample_n > 0) ? dit->second.sample.data() : nullptr;
    const int n_rows = sample ? dit->second.sample_n : config.n_part;
    outstream.open("./dumps/" + std::to_string(dump_counter) + ".txt");
    outstream << "# This file was dumped at t = " << current_time << std::endl
              << "# Column definitions:" << std::endl
              << "# Particle ID / Type / Smoothing length / Density / Pressure / Acceleration / Velocity / Position / Thermal energy" << std::endl
              << "# Aligned definition 'tags' for easier reading:" << std::endl
              << "# ID    TYPE     H          DENSITY  PRESS    ACCEL     VEL       POS       U" << std::endl;
    char line[256];
    for (int i = 0; i < n_rows; i++) {
        if (sample) {
            const sphb_particle &p = sample[i];
            std::snprintf(line, sizeof line, "%4d    %s    %3.5f    %3.3f    %3.3f    %+3.3f    %+3.3f    %+3.3f    %3.3f\n",
                          p.id, ParticleTypeNames[p.type], p.h, p.density, p.pressure, p.acc, p.vel, p.pos, p.u);
        } else {
            const Particle &p = p_arr[i];
            std::snprintf(line, sizeof line, "%4d    %s    %3.5f    %3.3f    %3.3f    %+3.3f    %+3.3f    %+3.3f    %3.3f\n",
                          p.id, ParticleTypeNames[p.type], p.h, p.density, p.pressure, p.acc, p.vel, p.pos, p.u);
        }
        outstream << line;
    }
    outstream.close();
    dump_counter++;
}
