// ghost_particles_sphb.cpp — setup_ghost_particles of the reference's sph/ghost_particles.hpp
// on top of libsphb (replaces sph/ghost_particles.cpp).  Like the reference it replaces the
// caller's array when the ghosts do not fit and rewrites config.n_part / config.n_ghost.
#include "ghost_particles.hpp"   // the reference's
#include "sphb_bridge.hpp"

// ghost slots the current array was allocated with (the reference's grow-only policy,
// ghost_particles.cpp:17,77-108)
static int allocated_ghost_slots = 0;

void setup_ghost_particles(ParticleArrayPtr &p_arr, Config &config) {
    using namespace sphb_adaptor;
    const int n_alive = config.n_part - config.n_ghost;
    Config alive_only = config;
    alive_only.n_part = n_alive;
    alive_only.n_ghost = 0;
    Bridge &b = Bridge::get();
    sphb_handle *h = b.handle(alive_only);
    b.invalidate();   // the functor image no longer describes the device state
    impose_fixed_h(reinterpret_cast<sphb_particle *>(p_arr.get()), n_alive);   // run-time fixed-h mode only
    int rc = sphb_upload(h, reinterpret_cast<const sphb_particle *>(p_arr.get()), n_alive, 0);
    int n_ghost = 0;
    if (!rc) rc = sphb_ghosts(h, &n_ghost);
    if (rc) fail(h, rc);
    if (n_ghost > allocated_ghost_slots) {
        std::cout << "[INFO] Reallocating particle array for " << n_ghost << " ghost particles" << std::endl;
        ParticleArrayPtr grown(new Particle[n_alive + n_ghost]);
        allocated_ghost_slots = n_ghost;
        p_arr.swap(grown);   // alive records are rewritten by the download below
    }
    int n_part = 0;
    rc = sphb_download(h, reinterpret_cast<sphb_particle *>(p_arr.get()), n_alive + n_ghost, &n_part, &n_ghost,
                       /*write_ids=*/0);   // Particle::id is const: leave the ids of the array alone
    if (rc) fail(h, rc);
    config.n_ghost = n_ghost;
    config.n_part = n_part;
}
