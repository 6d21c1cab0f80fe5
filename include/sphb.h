/*
 * include/sphb.h — C-ABI of libsphb.so, the B200-native replacement for the
 * per-timestep particle loop of SapiensAnatis/sph.
 *
 * Everything here is plain C: opaque handle, plain pointers and sizes, int
 * status returns (0 = ok), no exceptions cross the boundary, one host thread
 * per handle.  Each entry point names the reference interface it replaces
 * (paths relative to the reference repository).  The reference's API is a
 * per-particle C++ functor protocol (sph/calculators.hpp:19-127); a functor
 * cannot be accelerated call by call, so the ABI is batch-shaped and the C++
 * adaptors in sph_b200/adaptors/ keep the reference's names on top of it
 * (see INTEGRATION.md).
 */
#ifndef SPHB_H
#define SPHB_H

#ifdef __cplusplus
extern "C" {
#endif

/* Layout-identical to the reference's `Config` (sph/basictypes.hpp:21-31, 64 bytes). */
typedef struct sphb_config {
    int n_part;          /* alive + ghost, as the reference keeps it after ghost set-up */
    double mass;
    int pressure_calc;   /* 0 isothermal, 1 adiabatic (enum PressureCalc, basictypes.hpp:16-19) */
    double limit;
    double v_0;
    double h_factor;
    double t_i;
    int n_ghost;
} sphb_config;

/* Layout-identical to the reference's `Particle` (sph/basictypes.hpp:49-116, 88 bytes). */
typedef struct sphb_particle {
    int id;
    double mass;
    double pos;
    double vel;
    double acc;
    double h;
    double du_dt;
    double u;
    double density;
    double pressure;
    int type;            /* 0 alive, 1 ghost (enum ParticleType, basictypes.hpp:35-38) */
} sphb_particle;

enum { SPHB_ISOTHERMAL = 0, SPHB_ADIABATIC = 1 };
enum { SPHB_ORDER_REFERENCE = 0, SPHB_ORDER_JACOBI = 1 };

/* Run-time options.  The first two replace the reference's compile-time switches
 * USE_VARIABLE_H / CONSTANT_H (sph/define.hpp:27-29). */
typedef struct sphb_options {
    int variable_h;      /* 1: per-particle Newton/bisection root-find (default); 0: fixed h */
    double fixed_h;      /* h given to every particle by sphb_init_lattice when variable_h = 0 */
    int order;           /* SPHB_ORDER_REFERENCE: the reference's index-ordered in-place sweep
                            (sph_simulation.cpp:94-107), reproduced exactly; SPHB_ORDER_JACOBI:
                            one fused pass, every neighbour seen half-kicked */
    int max_sweeps;      /* cap on force sweeps per step in reference order (default 64) */
    int device;          /* CUDA device ordinal, -1 = current device */
    int pad_capacity;    /* max ghost (or halo) particles per side (default 4096) */
    int capacity;        /* max alive particles; 0 = config.n_part */
    int sort_every_step; /* 1: radix-sort every step; 0: only when the order check fails */
    double sweep_rel_tol; /* SPHB_ORDER_REFERENCE: 0 (default) iterates the ordered sweep to its BITWISE fixed point, i.e. the
                             reference's sequential result exactly; > 0 stops once no kicked v, u moves by more than this
                             relative amount (v against max(|v|, c_s)): bounded deviation, far fewer sweeps when every
                             particle's kick is non-zero */
    int sweep_mode;      /* SPHB_ORDER_REFERENCE: how the ordered sweep is computed.  Every choice gives the same bits (the
                            reference's sequential result); they differ in cost only.  SPHB_SWEEP_AUTO (default) picks per
                            step; SPHB_SWEEP_WORKLIST: first guess for all particles, then fixed-point refinement of the
                            tiles in reach of a changed particle (cheap while almost no kick is non-zero);
                            SPHB_SWEEP_SERIAL: one pass in which groups of lanes walk segments of the x-sorted array
                            left to right (cost independent of the state) */
    int sweep_group;     /* serial sweep: lanes per group, 8 or 16 (slots a particle looks back / ahead = lanes - 1); 0 = auto */
    int sweep_segment;   /* serial sweep: particles per segment; 0 = auto */
    int sweep_burn_in;   /* serial sweep: particles a segment starts early; 0 = auto (grows when its check fails) */
    int reserved[2];
} sphb_options;
enum { SPHB_SWEEP_AUTO = 0, SPHB_SWEEP_WORKLIST = 1, SPHB_SWEEP_SERIAL = 2 };

typedef struct sphb_counters {
    int n_alive;
    int n_ghost;
    long long steps;            /* steps taken by sphb_step since creation */
    long long sweeps_total;     /* force sweeps over all steps (>= steps) */
    int sweeps_last;            /* force sweeps of the most recent step */
    long long sorts;            /* x-order repairs performed: radix sorts + cheap transposition repairs */
    long long fallbacks;        /* bisection fallbacks (smoothing_length.cpp:246-254) */
    long long newton_hist[12];  /* most recent density pass: particles by Newton iteration count */
    int nbr_min, nbr_max;       /* most recent density pass: neighbours with |dx|/h < 2.5, incl. self */
    double nbr_mean;
    double h_max;
    long long launches;         /* kernels this handle has launched (CUB's sort kernels not counted) */
    double force_pairs_mean;    /* most recent step: (i,j) pairs per particle in the first force sweep, self included */
    double density_scans_mean;  /* most recent density pass: neighbour-window scans per particle (1 + Newton iterations, minus
                                   the final one where the last Newton step was below 1e-12 and the sums were reused) */
    int refine_tiles;           /* most recent step: tiles the ordered-sweep refinement took from its work lists */
    int refine_tile_evals;      /* ... and how often it evaluated them, in-tile iterations included */
    long long sweep_steps;      /* steps whose ordered sweep was done by the serial-sweep kernel */
    int sweep_group, sweep_segment, sweep_burn_in;   /* its geometry in the most recent step (0: that step used work lists) */
    int sweep_flagged;          /* most recent step: particles it handed to the work-list refinement */
    int sweep_verify_fail;      /* most recent step: burn-in values that differed from the left segment's final ones */
} sphb_counters;

/* status codes */
enum {
    SPHB_OK = 0,
    SPHB_ERR_INVALID = 1,          /* bad argument */
    SPHB_ERR_CUDA = 2,             /* CUDA runtime failure; text in sphb_last_error */
    SPHB_ERR_CAPACITY = 3,         /* more particles / ghosts than the handle was sized for */
    SPHB_ERR_NEGATIVE_H = 4,       /* reference: calculators.cpp:38-42 throws */
    SPHB_ERR_LOW_DENSITY = 5,      /* reference: calculators.cpp:174-186 throws */
    SPHB_ERR_GHOST_ZERO_H = 6,     /* reference: ghost_particles.cpp:37-43 throws */
    SPHB_ERR_ZERO_DERIVATIVE = 7,  /* GSL newton "derivative is zero" (aborts in the reference) */
    SPHB_ERR_NOT_FINITE = 8,       /* GSL "function value is not finite" (aborts) */
    SPHB_ERR_NO_BRACKET = 9,       /* GSL bisection "endpoints do not straddle" (aborts) */
    SPHB_ERR_NOT_CONVERGED = 10,   /* ordered sweep did not reach its fixed point in max_sweeps */
    SPHB_ERR_UNSORTED = 11,        /* internal: order invariant violated */
    SPHB_ERR_NO_DEVICE = 12,       /* no CUDA device / extension unusable: there is NO CPU fallback */
    SPHB_ERR_HALO = 13             /* x-slab run: the halo is thinner than the kernel support reaches */
};

typedef struct sphb_handle sphb_handle;

void sphb_default_options(sphb_options *opt);

/* Replaces: construction of the functors with (Config, ParticleArrayPtr)
 * (calculators.hpp:23-24,48-49,59-60,121-122) and SPHSimulation's constructor
 * (sph_simulation.hpp:22-24).  Allocates the device structure-of-arrays. */
int sphb_create(const sphb_config *cfg, const sphb_options *opt, sphb_handle **out);
void sphb_destroy(sphb_handle *h);
const char *sphb_last_error(const sphb_handle *h);

/* Run all device work of this handle on the given cudaStream_t (NULL = the handle's own). */
int sphb_set_stream(sphb_handle *h, void *cuda_stream);

/* Host array-of-structs <-> device structure-of-arrays.  Replaces Calculator::update
 * (calculators.hpp:31-34), i.e. "here is the (possibly reallocated) particle array".
 * aos holds n_part records, alive block first then n_ghost ghosts, in array-index order;
 * that order (which defines the reference's sweep, sph_simulation.cpp:94) is preserved. */
int sphb_upload(sphb_handle *h, const sphb_particle *aos, int n_part, int n_ghost);
/* write_ids = 0 leaves every record's `id` bytes untouched (it is `const int` on the
 * reference side); 1 writes the ids seen by the last upload (faster, one contiguous copy). */
int sphb_download(sphb_handle *h, sphb_particle *aos, int cap, int *n_part, int *n_ghost,
                  int write_ids);
/* Snapshot for dumps (replaces the full copy behind SPHSimulation::file_write, sph_simulation.cpp:110-139, when only
 * a sample of a large run is written): every `stride`-th alive record (array indices 0, stride, 2 stride, ...) followed
 * by ALL ghost records; only those cross the PCIe link.  n_alive_out / n_ghost_out receive the counts written. */
int sphb_download_strided(sphb_handle *h, sphb_particle *aos, int cap, int stride, int *n_alive_out,
                          int *n_ghost_out, int write_ids);
/* Same state as [n][9] doubles (mass,pos,vel,acc,h,du_dt,u,density,pressure) + type. */
int sphb_download_fields(sphb_handle *h, double *fields, int *type, int cap, int *n_part,
                         int *n_ghost);
/* Per-particle outputs of the most recent density pass, array-index order. */
int sphb_download_counts(sphb_handle *h, int *newton_iters, int *neighbours, int cap);

/* Replaces init_particles (setup.cpp:143-228): lattice, +-v0 / +-c_s streams, u0, h0,
 * adiabatic pre-pass, ghosts, t = 0 density / acceleration / du_dt — all on the device.
 * The GHOST records' acc and du_dt are 0 after this call.  In the reference they hold two values nothing ever reads:
 * the pre-pass acceleration (computed from a `pressure` member that is still uninitialised there, setup.cpp:192-202)
 * and a du_dt evaluated at the mirrored position (setup.cpp:223-226).  Every other ghost field equals the reference's;
 * from the first step on ghosts are copies of their source and acc / du_dt are copied with them. */
int sphb_init_lattice(sphb_handle *h);

/* No counterpart in the reference (its only start state is the lattice): a synthetic DEVELOPED state of the same
 * colliding adiabatic streams — the exact solution at time t_star (gas at rest at the walls, rarefaction fans, the
 * undisturbed streams, two outward shocks around shocked gas at rest) sampled with equal-mass particles, then ghosts,
 * density / root-find (ghosts included) and forces as in init_particles.  For benchmarks and tests on a state in which
 * every particle is being accelerated; t_star < ~0.55 with the default limit (the fans must not have met the shocks). */
int sphb_init_developed(sphb_handle *h, double t_star);

/* Replaces DensityCalculator::operator() (calculators.cpp:33-55) and rootfind_h /
 * calc_density / calc_omega (smoothing_length.cpp:64-91,191-259) for every alive particle
 * (and every ghost when include_ghosts != 0 — the t = 0 case, setup.cpp:218-220). */
int sphb_density(sphb_handle *h, int include_ghosts);
/* Replaces AccelerationCalculator::operator() + EnergyCalculator::operator()
 * (calculators.cpp:63-129,192-213) for every alive particle, no kick
 * (the order-independent t = 0 pass, setup.cpp:223-226). */
int sphb_forces(sphb_handle *h);
/* Replaces calc_omega (smoothing_length.cpp:64-77) for every particle, ghosts included:
 * omega[i] = 1 + (h_i / rho_i) sum_j m_j dW/dh at the particle's STORED h and density and the
 * current positions (1 for fixed h); array-index order.  Nothing of the state changes. */
int sphb_omega(sphb_handle *h, double *omega, int cap);
/* Replaces setup_ghost_particles (ghost_particles.cpp:19-115). */
int sphb_ghosts(sphb_handle *h, int *n_ghost);
/* Replaces the first loop of SPHSimulation::step_forward (sph_simulation.cpp:59-73). */
int sphb_kick_drift(sphb_handle *h);
/* Replaces SPHSimulation::step_forward (sph_simulation.cpp:55-108), nsteps times; the
 * state stays in HBM between steps. */
int sphb_step(sphb_handle *h, int nsteps);

int sphb_get_counters(sphb_handle *h, sphb_counters *out);
/* No counterpart in the reference (its author lists conservation checks as missing, readme.md:23): totals over
 * the alive particles of this handle, out4 = { sum m, sum m v, sum m v^2 / 2, sum m u }, summed in a fixed order. */
int sphb_diagnostics(sphb_handle *h, double *out4);
int sphb_get_config(const sphb_handle *h, sphb_config *out);

/* Per-phase device time, CUDA events on the handle's stream.  Phases:
 * 0 kick_drift, 1 sort, 2 ghosts, 3 density, 4 force sweep 1, 5 ordered-sweep refinement. */
enum { SPHB_PHASES = 6 };
int sphb_set_profiling(sphb_handle *h, int on);
/* ms[p], launches[p] accumulated since the last reset; reset != 0 clears them */
int sphb_get_phase_times(sphb_handle *h, double *ms, long long *launches, int reset);

/* ---- x-slab decomposition across GPUs -------------------------------------------------------
 * No counterpart in the reference (it is single-threaded); these entry points run
 * SPHSimulation::step_forward (sph_simulation.cpp:55-108) and init_particles (setup.cpp:143-228)
 * phase by phase on ONE slab of an x-sorted decomposition, so that the caller can exchange halo
 * messages with the neighbouring ranks between the phases.  One handle per rank.  The library packs
 * and unpacks messages in caller-owned DEVICE buffers of sphb_slab_message_doubles() doubles; moving
 * them (NCCL send/recv, or a device copy) is the caller's job.  side: -1 = left neighbour, +1 = right.
 * Step: step_begin | exchange BASE | density | exchange RECORDS | force_first, refine |
 *       { exchange ITERATE | refine } a fixed number of rounds | exchange ITERATE | poll (the step's one
 *       host synchronisation) | ranks agree on MAX(flags): error -> raise; order broken -> resort and
 *       repeat from 'exchange BASE'; halo changed -> more rounds; else agree (halo width), step_end.
 * With the serial sweep (sphb_options.sweep_mode) force_first walks the slab once, the left halo included, and the
 * ITERATE exchanges only confirm what it computed for the halo particles.                            */
enum { SPHB_MSG_BASE = 0, SPHB_MSG_RECORDS = 1, SPHB_MSG_ITERATE = 2 };
/* messages: BASE 8 rows, RECORDS 9 rows, ITERATE 2 rows of halo_count doubles, plus a header */
/* this handle holds global particle indices [index_offset, index_offset + n_alive) of n_global;
 * halo_count particles travel per edge (must exceed the kernel reach in particles; too thin a halo is
 * detected on the device and reported as SPHB_ERR_HALO).  A particle that overtakes its neighbour across a
 * slab edge changes owner (both ranks merge-split the 2 x halo_count particles around the edge; counts stay
 * equal), so a slab's particles are no longer a contiguous index range: in an x-slab run sphb_upload /
 * sphb_download move the records in x ORDER and `id` carries the particle's GLOBAL index (the caller sets
 * it on upload and reads it on download); sphb_init_lattice-style set-up needs no upload. */
int sphb_slab_configure(sphb_handle *h, int rank, int world, int halo_count, long long index_offset,
                        long long n_global);
int sphb_slab_message_doubles(const sphb_handle *h, int what);
int sphb_slab_pack(sphb_handle *h, int what, int side, double *d_buf);
int sphb_slab_unpack(sphb_handle *h, int what, int side, const double *d_buf);
int sphb_slab_step_begin(sphb_handle *h);   /* kick-drift, order check (+ sort), wall ghosts */
int sphb_slab_density(sphb_handle *h);      /* density + root-find + omega + records of own particles */
int sphb_slab_force_first(sphb_handle *h);  /* first force sweep + second kick */
int sphb_slab_refine(sphb_handle *h);       /* ordered-sweep refinement until the local work list is empty */
int sphb_slab_poll(sphb_handle *h, int *flags6);  /* {device error code, x-order broken, halo iterate changed, h max as float bits,
                                                      thinner halo of this rank (coded so that MAX = thinnest), work lists busy} */
/* with the element-wise MAX over ranks of the polled flags: does the widest kernel support of the run fit inside its thinnest
 * halo (every rank gets the same answer, so all leave the step together), and the sweep strategy of the next step */
int sphb_slab_agree(sphb_handle *h, const int *flags6_max);
int sphb_slab_resort(sphb_handle *h);       /* after a poll saw 'order broken' on some rank */
int sphb_slab_raise(sphb_handle *h);        /* the polled error code as status + message */
int sphb_slab_step_end(sphb_handle *h);
/* t = 0 in phases: init_lattice | [adiabatic: exchange BASE | sphb_density(h,0) | prepass] | sphb_ghosts |
 * exchange BASE | sphb_density(h,1) | records | exchange RECORDS | forces */
int sphb_slab_init_lattice(sphb_handle *h);
int sphb_slab_prepass(sphb_handle *h);
int sphb_slab_records(sphb_handle *h);
int sphb_slab_forces(sphb_handle *h);

/* The same protocol driven inside the library: halo messages go to the direct neighbours with
 * grouped ncclSend/ncclRecv on the handle's stream (libnccl is opened at run time).  Rank 0 makes
 * an id with sphb_nccl_unique_id, the host program distributes its 128 bytes (MPI_Bcast,
 * torch.distributed, a file ...), every rank calls sphb_slab_configure and sphb_slab_comm_init, then
 * all ranks make the same sphb_slab_init_particles / sphb_slab_step calls. */
int sphb_nccl_unique_id(void *out128);
int sphb_slab_comm_init(sphb_handle *h, const void *unique_id128);
int sphb_slab_init_particles(sphb_handle *h);   /* init_particles, setup.cpp:143-228, across the slabs */
int sphb_slab_init_developed(sphb_handle *h, double t_star);   /* sphb_init_developed across the slabs */
int sphb_slab_init_developed_fill(sphb_handle *h, double t_star);  /* its first phase only: this handle's particles, no ghosts */
int sphb_slab_step(sphb_handle *h, int nsteps); /* step_forward, sph_simulation.cpp:55-108, across the slabs */
int sphb_slab_rounds_last(const sphb_handle *h);/* ITERATE exchanges the last step needed */

/* No counterpart in the reference.  The same density / root-find / omega pass as sphb_density,
 * for callers that cannot allocate device memory when they need it (the whole device already
 * belongs to their own arena allocator, or they are inside a section where cudaMalloc's
 * device-wide synchronisation is not acceptable).  No handle: d_x (ascending), d_h, d_rho are
 * device arrays of n particles of equal `mass` with no ghosts; results go to d_h_out, d_rho_out
 * (device, n each; for variable_h = 0 only d_rho_out changes meaningfully, d_h_out = d_h).
 * Scratch is a fixed device array inside the library, so n <= SPHB_STATIC_MAX_PARTICLES
 * (larger n returns SPHB_ERR_CAPACITY and launches nothing), every call shares it, and calls
 * that overlap on the GPU race: serialise them.  The first call on a device resolves the
 * array's address; make it before any stream capture.  Synchronises `cuda_stream`
 * (NULL = the default stream) and returns 0 or the first device-side error code; the text
 * is read with sphb_last_error(NULL).  Calls no cudaMalloc / cudaFree. */
#define SPHB_STATIC_MAX_PARTICLES (1 << 20)
int sphb_density_static(const double *d_x, const double *d_h, const double *d_rho, int n, double mass,
                        double h_factor, double limit, int variable_h, double *d_h_out, double *d_rho_out,
                        void *cuda_stream);

/* library identity: returns the compiled-for architecture string, e.g. "sm_100a" */
const char *sphb_build_arch(void);

#ifdef __cplusplus
}
#endif
#endif
