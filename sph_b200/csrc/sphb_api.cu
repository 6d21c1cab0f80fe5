// sphb_api.cu — host side of libsphb.so: the C-ABI of include/sphb.h over the kernels
// in sphb_kernels.cuh.  No CPU fallback exists: without a usable CUDA device every entry
// point fails with SPHB_ERR_NO_DEVICE / SPHB_ERR_CUDA.
#include <cub/device/device_radix_sort.cuh>
#include <dlfcn.h>
#include <nccl.h>   // types only: the library is opened at run time (sphb_slab_comm_init)

#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdio>
#include <cstring>
#include <string>
#include <utility>
#include <vector>

#include "../../include/sphb.h"
#include "sphb_kernels.cuh"

using namespace sphb;

static_assert(sizeof(sphb_particle) == 88, "sphb_particle must match the reference Particle");
static_assert(sizeof(sphb_config) == 64, "sphb_config must match the reference Config");

struct sphb_handle {
    sphb_config cfg;
    sphb_options opt;
    int device = 0;
    int num_sms = 0;
    cudaStream_t stream = nullptr, own_stream = nullptr;
    int pad = 0, cap_alive = 0, total = 0, cand_cap = 0;
    int na = 0;
    // state (current): positions, velocity, energy, derivatives, h, rho, stored pressure, mass
    double *x = nullptr, *v = nullptr, *u = nullptr, *a = nullptr, *du = nullptr, *h = nullptr,
           *rho = nullptr, *P = nullptr, *m = nullptr;
    // per-step work arrays
    double *xalt = nullptr, *vh = nullptr, *uh = nullptr, *vn = nullptr, *un = nullptr, *hn = nullptr,
           *rhon = nullptr, *omn = nullptr, *oms = nullptr, *Pn = nullptr, *tmp = nullptr, *an = nullptr,
           *dun = nullptr;   // acc, du_dt of the step in flight (swapped in by step_finish)
    unsigned *idx = nullptr, *idx_alt = nullptr;
    int *slot_id = nullptr, *cnt = nullptr, *dirty = nullptr, *cand = nullptr, *gorder = nullptr,
        *perm = nullptr, *iota = nullptr, *wl0 = nullptr, *wl1 = nullptr, *wl2 = nullptr;
    int refine_peak = 0;      // largest refinement work list of the previous step
    Rec2 *rec = nullptr;      // force-pass records written by the density pass
    // serial sweep (k_force_sweep): per-tile h maximum of the density pass, burn-in results kept for the check
    unsigned *thm = nullptr;
    double *scr = nullptr;
    size_t scr_cap = 0;       // doubles
    bool sweep_on = true;     // strategy of the next step in auto mode: serial sweep (true) or work lists (false)
    bool sweep_used = false;  // ... and what the step in flight uses
    int sw_G = 8, sw_L = 0, sw_B = 0, sw_nseg = 0;
    int sweep_blocks[2] = {0, 0};   // resident CTAs per SM of the G = 8 / G = 16 kernels (0 = not asked yet)
    long long sweep_steps = 0;
    bool agreed_busy = true;      // x-slab runs: MAX over ranks of sweep_busy() for the step in flight
    bool sweep_verify_pending = false;
    int sweep_flagged_last = 0, sweep_verify_fail_last = 0, sweep_burn_in_last = 0;
    double *gsum = nullptr;   // own-h gradient sums written by the density pass
    // ghost staging (2*pad each)
    double *gx = nullptr, *gv = nullptr, *gu = nullptr, *ga = nullptr, *gdu = nullptr, *gh = nullptr,
           *grho = nullptr, *gP = nullptr, *gm = nullptr;
    unsigned *gkey = nullptr;
    Ctl *d_ctl = nullptr;
    Ctl *h_ctl = nullptr;  // pinned
    void *cub_tmp = nullptr;
    size_t cub_bytes = 0;
    AosRec *d_aos = nullptr;
    size_t aos_cap = 0;
    int *d_newton = nullptr, *d_nbr = nullptr;  // index-ordered count outputs (lazy)
    bool ghost_in_half = false;   // ghosts' v,u live in the pads of vh/uh (after a step)
    // x-slab decomposition (sphb_slab_*): this handle owns global indices [idx_base, idx_base + na)
    bool slab = false, halo_l = false, halo_r = false;
    int rank = 0, world = 1, K = 0;
    long long idx_base = 0, n_global = 0;
    // native exchange (sphb_slab_comm_init): NCCL communicator, message buffers [what][side], agreed flags
    ncclComm_t comm = nullptr;
    double *msg_send[3][2] = {}, *msg_recv[3][2] = {};
    int *d_flags = nullptr, *h_flags = nullptr;
    int fixed_rounds = 1, rounds_last = 0;
    cudaStream_t comm_stream = nullptr;          // BASE / RECORDS exchanges run here, beside the interior tiles
    cudaEvent_t ev_kick = nullptr, ev_base = nullptr, ev_edge = nullptr, ev_rec = nullptr;
    int part = 0;                                // 0: all tiles, 1: interior tiles only, 2: edge tiles only
    bool uniform_mass = true;
    double m0 = 0.0;
    bool have_counts = false;
    long long steps = 0, sweeps_total = 0, sorts = 0, swap_repairs = 0;
    int resorts_this_step = 0;
    unsigned long long fallbacks_committed = 0;   // bisection fallbacks of finished steps (a repeated attempt restarts from it)
    int sweeps_last = 0;
    int epoch = 8;
    int refine_grid = 0;
    long long launches = 0;
    bool profiling = false;
    cudaEvent_t ev[SPHB_PHASES + 1] = {};
    double phase_ms[SPHB_PHASES] = {};
    long long phase_n[SPHB_PHASES] = {};
    bool phase_hit[SPHB_PHASES] = {};
    std::string err;
};

namespace {

constexpr int kSwapPasses = 4;   // odd-even transposition rounds of the cheap order repair

#define CK(call)                                                                              \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess) {                                                              \
            char b_[512];                                                                     \
            snprintf(b_, sizeof b_, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_),   \
                     __FILE__, __LINE__);                                                     \
            H->err = b_;                                                                      \
            return SPHB_ERR_CUDA;                                                             \
        }                                                                                     \
    } while (0)

#define TRY(expr)                 \
    do {                          \
        int rc_ = (expr);         \
        if (rc_ != SPHB_OK) return rc_; \
    } while (0)

inline int nblk(int n, int b) { return (n + b - 1) / b; }

const char *err_text(int code) {
    switch (code) {
        case SPHB_ERR_CAPACITY: return "more ghost particles than pad_capacity allows";
        case SPHB_ERR_NEGATIVE_H: return "Root-finding returned negative smoothing length";
        case SPHB_ERR_LOW_DENSITY: return "Particle had density less than epsilon!";
        case SPHB_ERR_GHOST_ZERO_H: return "Ghost particle initialization: Particle had zero smoothing length!";
        case SPHB_ERR_ZERO_DERIVATIVE: return "newton: derivative is zero";
        case SPHB_ERR_NOT_FINITE: return "root-finding: function value is not finite";
        case SPHB_ERR_NO_BRACKET: return "bisection: endpoints do not straddle y=0";
        case SPHB_ERR_NOT_CONVERGED: return "ordered sweep did not converge within max_sweeps";
        case SPHB_ERR_UNSORTED: return "x-order invariant violated";
        case SPHB_ERR_HALO: return "x-slab halo thinner than the kernel support (raise halo_count)";
        default: return "error";
    }
}

constexpr size_t kErrBytes = 2 * sizeof(int) + sizeof(unsigned long long);   // err_code, err_unused, err_first
static_assert(offsetof(Ctl, err_first) == offsetof(Ctl, err_code) + 2 * sizeof(int), "the error words are cleared with one memset");

// The device error word -> status + message.  Reported is the LOWEST sorted slot that raised and ITS code: with the
// array order equal to the x order that is the particle the reference's sequential loops fail at first.
int decode_error(const Ctl &c, const unsigned *d_idx, char *msg, size_t cap) {
    const unsigned long long packed = ~c.err_first;
    const int code = (int)(packed & 0xff);
    const int slot = (int)(packed >> 8) - 1;
    if (slot < 0) { snprintf(msg, cap, "%s", err_text(code)); return code; }
    unsigned pidx = 0;
    if (d_idx && cudaMemcpy(&pidx, d_idx + slot, sizeof pidx, cudaMemcpyDeviceToHost) == cudaSuccess) {
        if (pidx & kGhostBit) snprintf(msg, cap, "%s (ghost of particle %u, sorted slot %d)", err_text(code), pidx & ~kGhostBit, slot);
        else snprintf(msg, cap, "%s (particle %u, sorted slot %d)", err_text(code), pidx, slot);
    } else {
        snprintf(msg, cap, "%s (particle %d)", err_text(code), slot);
    }
    return code;
}

template <class T>
int dalloc(sphb_handle *H, T **p, size_t n) {
    CK(cudaMalloc((void **)p, sizeof(T) * std::max<size_t>(n, 1)));
    CK(cudaMemsetAsync(*p, 0, sizeof(T) * std::max<size_t>(n, 1), H->stream));
    return SPHB_OK;
}

// Copy the control block to the host (one small D2H + stream sync) and turn a device
// error word into a status.
int read_ctl(sphb_handle *H) {
    CK(cudaMemcpyAsync(H->h_ctl, H->d_ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, H->stream));
    CK(cudaStreamSynchronize(H->stream));
    if (H->h_ctl->err_code != 0) {
        char b[256];
        const int code = decode_error(*H->h_ctl, H->idx, b, sizeof b);
        H->err = b;
        // clear so that the handle stays usable after the caller has seen the error
        CK(cudaMemsetAsync(&H->d_ctl->err_code, 0, kErrBytes, H->stream));
        return code;
    }
    return SPHB_OK;
}

// ghosts made at the walls (pads that hold a neighbouring slab's particles do not count)
int wall_ghosts(const sphb_handle *H) {
    return (H->halo_l ? 0 : H->h_ctl->GL) + (H->halo_r ? 0 : H->h_ctl->GR);
}

void phase_begin(sphb_handle *H) {
    if (H->profiling) cudaEventRecord(H->ev[0], H->stream);
}
// records the end of phase p; phases of one step are recorded back to back
void phase_end(sphb_handle *H, int p) {
    if (!H->profiling) return;
    cudaEventRecord(H->ev[p + 1], H->stream);
    H->phase_hit[p] = true;
}
void phase_collect(sphb_handle *H) {
    if (!H->profiling) return;
    cudaEvent_t prev = H->ev[0];
    for (int p = 0; p < SPHB_PHASES; p++) {
        if (!H->phase_hit[p]) continue;
        float ms = 0.f;
        cudaEventSynchronize(H->ev[p + 1]);
        if (cudaEventElapsedTime(&ms, prev, H->ev[p + 1]) == cudaSuccess) {
            H->phase_ms[p] += ms;
            H->phase_n[p]++;
        }
        prev = H->ev[p + 1];
        H->phase_hit[p] = false;
    }
}

int launch_begin(sphb_handle *H, int clear_cand) {
    k_begin<<<1, 1, 0, H->stream>>>(H->d_ctl, clear_cand);
    H->launches++;
    CK(cudaGetLastError());
    return SPHB_OK;
}

// kick + drift + ghost selection + order check (do_kick = 1), or selection + check only
int launch_kick(sphb_handle *H, int do_kick) {
    KickArgs A;
    A.x = H->x; A.xo = H->xalt; A.v = H->v; A.u = H->u; A.a = H->a; A.du = H->du; A.h = H->h;
    A.vh = H->vh; A.uh = H->uh; A.cand = H->cand; A.ctl = H->d_ctl;
    A.pad = H->pad; A.na = H->na; A.cand_cap = H->cand_cap;
    A.dt = H->cfg.t_i; A.half_dt = H->cfg.t_i / 2; A.limit = H->cfg.limit;
    A.do_kick = do_kick;
    if (H->na > 0) {
        k_kick_drift<<<nblk(H->na, kTile), kTile, 0, H->stream>>>(A);
        H->launches++;
        CK(cudaGetLastError());
    }
    if (do_kick) std::swap(H->x, H->xalt);
    return SPHB_OK;
}

int gather_f64(sphb_handle *H, double *&arr) {
    k_gather_f64<<<nblk(H->na, 256), 256, 0, H->stream>>>(arr, H->tmp, H->perm, H->pad, H->na);
    H->launches++;
    CK(cudaGetLastError());
    std::swap(arr, H->tmp);
    return SPHB_OK;
}

// Radix-sort the alive block by x (CUB DeviceRadixSort on the FP64 keys with the slot
// number as payload) and permute the listed arrays.  `full` = every state array (upload
// path); otherwise only what is live between the drift and the density pass.
int sort_alive(sphb_handle *H, bool full) {
    if (H->na <= 1) return SPHB_OK;
    k_iota<<<nblk(H->na, 256), 256, 0, H->stream>>>(H->iota, H->na);
    H->launches++;
    CK(cudaGetLastError());
    size_t bytes = H->cub_bytes;
    CK(cub::DeviceRadixSort::SortPairs(H->cub_tmp, bytes, H->x + H->pad, H->tmp + H->pad, H->iota, H->perm,
                                       H->na, 0, 64, H->stream));
    std::swap(H->x, H->tmp);   // sorted keys are the new positions
    if (full) {
        TRY(gather_f64(H, H->v)); TRY(gather_f64(H, H->u)); TRY(gather_f64(H, H->a)); TRY(gather_f64(H, H->du));
    } else {
        TRY(gather_f64(H, H->vh)); TRY(gather_f64(H, H->uh));
        // a, du of the alive block are rewritten by the force pass, but the ghost copies made
        // right after this sort take them from their source's slot (ghost_particles.cpp:53,111)
        TRY(gather_f64(H, H->a)); TRY(gather_f64(H, H->du));
    }
    TRY(gather_f64(H, H->h)); TRY(gather_f64(H, H->rho)); TRY(gather_f64(H, H->P)); TRY(gather_f64(H, H->m));
    k_gather_u32<<<nblk(H->na, 256), 256, 0, H->stream>>>(H->idx, H->idx_alt, H->perm, H->pad, H->na);
    H->launches++;
    CK(cudaGetLastError());
    std::swap(H->idx, H->idx_alt);
    H->sorts++;
    return SPHB_OK;
}

// Place ghosts into the pads. from_alive: mirror the selected candidates out of (vs,us);
// otherwise the staging arrays were filled by the upload.  Ghost v,u go to the pads of (vd,ud).
int launch_ghost_build(sphb_handle *H, int from_alive, int nstaged, double *vs, double *us, double *vd, double *ud,
                       bool into_next = false) {
    GhostArgs G;
    G.ad = into_next ? H->an : H->a; G.dud = into_next ? H->dun : H->du;
    G.x = H->x; G.vd = vd; G.ud = ud; G.a = H->a; G.du = H->du; G.h = H->h; G.rho = H->rho; G.P = H->P; G.m = H->m;
    G.vs = vs; G.us = us; G.idx = H->idx; G.cand = H->cand;
    G.gx = H->gx; G.gv = H->gv; G.gu = H->gu; G.ga = H->ga; G.gdu = H->gdu; G.gh = H->gh; G.grho = H->grho;
    G.gP = H->gP; G.gm = H->gm; G.gkey = H->gkey; G.gorder = H->gorder;
    G.ctl = H->d_ctl; G.pad = H->pad; G.na = H->na; G.cand_cap = H->cand_cap; G.limit = H->cfg.limit;
    G.from_alive = from_alive; G.nstaged = nstaged;
    k_ghost_build<<<1, 1024, 0, H->stream>>>(G);
    H->launches++;
    CK(cudaGetLastError());
    return SPHB_OK;
}

int launch_density(sphb_handle *H, int root_alive, int root_ghosts, int check_old, double *Pn,
                   bool want_omega = false) {
    DensityArgs A;
    A.x = H->x; A.m = H->m; A.h = H->h; A.rho = H->rho; A.P = H->P;
    A.hn = H->hn; A.rhon = H->rhon; A.Pn = Pn;
    A.omn = want_omega ? H->omn : nullptr; A.oms = want_omega ? H->oms : nullptr;
    A.rec = H->rec; A.gsum = H->gsum;
    A.cnt = H->cnt; A.ctl = H->d_ctl; A.pad = H->pad; A.na = H->na;
    A.limit = H->cfg.limit; A.hfac = H->cfg.h_factor; A.m0 = H->m0;
    A.root_ghosts = root_ghosts; A.root_alive = root_alive; A.check_old = check_old;
    A.thm = H->thm;
    const int ntot = nblk(H->na + 2 * H->pad, kTile);
    // tile ranges [begin, end) to launch: everything, or (x-slab overlap) the interior / the two edges;
    // edge = every tile that can see a pad: the pads themselves plus K + one tile of own particles
    int ranges[2][2] = {{0, ntot}, {0, 0}};
    if (H->part) {
        const int e = H->K + kTile;
        const int tl = std::min(ntot, nblk(H->pad + e, kTile));
        const int tr = std::max(tl, (H->pad + H->na - e) / kTile);
        if (H->part == 1) { ranges[0][0] = tl; ranges[0][1] = tr; }
        else { ranges[0][0] = 0; ranges[0][1] = tl; ranges[1][0] = tr; ranges[1][1] = ntot; }
    }
    const bool varh = H->opt.variable_h != 0;
#define SPHB_DENS(V, U)                                                              \
    do {                                                                             \
        if (H->slab) k_density<V, U, true><<<grid, kTile, 0, H->stream>>>(A);        \
        else k_density<V, U, false><<<grid, kTile, 0, H->stream>>>(A);               \
    } while (0)
    for (auto &rg : ranges) {
        const int grid = rg[1] - rg[0];
        if (grid <= 0) continue;
        A.tile_base = rg[0];
        if (varh && H->uniform_mass) SPHB_DENS(true, true);
        else if (varh) SPHB_DENS(true, false);
        else if (H->uniform_mass) SPHB_DENS(false, true);
        else SPHB_DENS(false, false);
        H->launches++;
        CK(cudaGetLastError());
    }
#undef SPHB_DENS
    if (root_alive || root_ghosts) H->have_counts = true;
    return SPHB_OK;
}

template <int SEL>
int launch_first_sel(sphb_handle *H, SweepArgs A) {
    const int ntiles = nblk(H->na, kTile);
    int ranges[2][2] = {{0, ntiles}, {0, 0}};
    if (H->part) {
        const int te = std::min(ntiles, nblk(H->K + kTile, kTile));
        const int tr = std::max(te, ntiles - te);
        if (H->part == 1) { ranges[0][0] = te; ranges[0][1] = tr; }
        else { ranges[0][0] = 0; ranges[0][1] = te; ranges[1][0] = tr; ranges[1][1] = ntiles; }
    }
    const bool adia = H->cfg.pressure_calc == SPHB_ADIABATIC;
#define SPHB_FIRST(AD, U)                                                                        \
    do {                                                                                         \
        if (H->part) k_force_first<SEL, AD, U, true><<<grid, kTile, 0, H->stream>>>(A);          \
        else k_force_first<SEL, AD, U, false><<<grid, kTile, 0, H->stream>>>(A);                 \
    } while (0)
    for (auto &rg : ranges) {
        const int grid = rg[1] - rg[0];
        if (grid <= 0) continue;
        A.tile_base = rg[0];
        if (adia && H->uniform_mass) SPHB_FIRST(true, true);
        else if (adia) SPHB_FIRST(true, false);
        else if (H->uniform_mass) SPHB_FIRST(false, true);
        else SPHB_FIRST(false, false);
        H->launches++;
        CK(cudaGetLastError());
    }
#undef SPHB_FIRST
    return SPHB_OK;
}

int launch_refine(sphb_handle *H, SweepArgs A) {
    const bool adia = H->cfg.pressure_calc == SPHB_ADIABATIC;
    void *fn;
    if (adia && H->uniform_mass) fn = (void *)k_force_refine<true, true>;
    else if (adia) fn = (void *)k_force_refine<true, false>;
    else if (H->uniform_mass) fn = (void *)k_force_refine<false, true>;
    else fn = (void *)k_force_refine<false, false>;
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, kTile, 0));
    if (occ < 1) { H->err = "refinement kernel cannot be made resident"; return SPHB_ERR_CUDA; }
    const int ntiles = nblk(H->na, kTile);
    // a grid-wide barrier costs more the more CTAs take part: the few tiles around a front get one CTA
    // per SM; a step whose predecessor queued many tiles gets every CTA that can be resident
    const int want = H->refine_peak > 2 * H->num_sms ? occ * H->num_sms : H->num_sms;
    int grid = std::min(ntiles, want);
    if (grid < 1) grid = 1;
    int max_sweeps = H->opt.max_sweeps;
    void *args[] = {(void *)&A, (void *)&max_sweeps};
    CK(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(kTile), args, 0, H->stream));
    H->launches++;
    return SPHB_OK;
}

SweepArgs step_sweep_args(sphb_handle *H);

template <int G>
int sweep_occupancy(sphb_handle *H, bool adia, int *blocks) {
    const size_t smem = (size_t)(kSweepThreads / G) * sizeof(SweepGrp<G>);
    const void *fn = adia ? (const void *)k_force_sweep<true, G> : (const void *)k_force_sweep<false, G>;
    CK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks, fn, kSweepThreads, smem));
    if (*blocks < 1) { H->err = "serial sweep kernel cannot be made resident"; return SPHB_ERR_CUDA; }
    return SPHB_OK;
}

// Does the step in flight use the serial sweep, and with what geometry?  Group width G: 8 lanes look 7 slots
// back and ahead (variable h keeps ~5 neighbours per side), 16 lanes 15 (fixed h of 5 spacings: 12 per side).
// Burn-in B: the left boundary's influence falls by an order of magnitude per neighbour hop, 16 decades need
// about 16 hops of up to (neighbours per side) slots.  Segment length L: long enough to make the burn-in cheap,
// short enough to give every resident group a segment, and a whole number of waves.
int sweep_plan(sphb_handle *H) {
    H->sweep_used = false;
    if (H->opt.order != SPHB_ORDER_REFERENCE || !H->uniform_mass || H->opt.sweep_mode == SPHB_SWEEP_WORKLIST) return SPHB_OK;
    if (H->opt.sweep_mode == SPHB_SWEEP_AUTO && !H->sweep_on) return SPHB_OK;
    if (H->na < 2) return SPHB_OK;
    const bool adia = H->cfg.pressure_calc == SPHB_ADIABATIC;
    const int G = H->opt.sweep_group == 16 ? 16 : (H->opt.sweep_group == 8 ? 8 : (H->opt.variable_h ? 8 : 16));
    int &blocks = H->sweep_blocks[G == 16 ? 1 : 0];
    if (!blocks) { if (G == 8) TRY(sweep_occupancy<8>(H, adia, &blocks)); else TRY(sweep_occupancy<16>(H, adia, &blocks)); }
    int B = H->opt.sweep_burn_in > 0 ? H->opt.sweep_burn_in : (H->sweep_burn_in_last ? H->sweep_burn_in_last : (G == 8 ? 64 : 192));
    B = std::max(G, (B + G - 1) / G * G);
    int L = H->opt.sweep_segment;
    if (L <= 0) {
        const long long resident = (long long)H->num_sms * blocks * (kSweepThreads / G);
        const long long lmax = 2048;
        const long long waves = std::max(1ll, (H->na + resident * lmax - 1) / (resident * lmax));
        L = (int)((H->na + resident * waves - 1) / (resident * waves));
        // several waves: keep the burn-in below a quarter of the work.  A single wave that does not even fill the
        // machine is latency-bound (a group's time is proportional to L + B): shorter segments, down to the burn-in
        // length, finish sooner although they do more work in total.
        L = std::max(L, waves > 1 ? 4 * B : B);
    }
    L = std::max(B, (L + G - 1) / G * G);
    H->sw_G = G; H->sw_L = L; H->sw_B = B; H->sw_nseg = nblk(H->na, L);
    if ((size_t)H->sw_nseg * 2 * (G - 1) > H->scr_cap) return SPHB_OK;   // cannot happen with the sizes of sphb_create
    H->sweep_used = true;
    return SPHB_OK;
}

int launch_sweep(sphb_handle *H) {
    SerialArgs K;
    K.S = step_sweep_args(H);
    K.thm = H->thm; K.nthm = nblk(H->na + 2 * H->pad, kTile);
    K.scr = H->scr; K.L = H->sw_L; K.B = H->sw_B; K.nseg = H->sw_nseg;
    const bool adia = H->cfg.pressure_calc == SPHB_ADIABATIC;
    if (H->sw_G == 8) {
        const int gpb = kSweepThreads / 8;
        const size_t smem = (size_t)gpb * sizeof(SweepGrp<8>);
        if (adia) k_force_sweep<true, 8><<<nblk(K.nseg, gpb), kSweepThreads, smem, H->stream>>>(K);
        else k_force_sweep<false, 8><<<nblk(K.nseg, gpb), kSweepThreads, smem, H->stream>>>(K);
    } else {
        const int gpb = kSweepThreads / 16;
        const size_t smem = (size_t)gpb * sizeof(SweepGrp<16>);
        if (adia) k_force_sweep<true, 16><<<nblk(K.nseg, gpb), kSweepThreads, smem, H->stream>>>(K);
        else k_force_sweep<false, 16><<<nblk(K.nseg, gpb), kSweepThreads, smem, H->stream>>>(K);
    }
    H->launches++;
    H->sweep_verify_pending = true;
    CK(cudaGetLastError());
    return SPHB_OK;
}

int launch_sweep_verify(sphb_handle *H) {
    if (H->sw_nseg < 2) return SPHB_OK;
    SerialArgs K;
    K.S = step_sweep_args(H);
    K.thm = H->thm; K.nthm = 0; K.scr = H->scr; K.L = H->sw_L; K.B = H->sw_B; K.nseg = H->sw_nseg;
    const int R = H->sw_G - 1;
    k_sweep_verify<<<nblk((H->sw_nseg - 1) * R, 256), 256, 0, H->stream>>>(K, R);
    H->launches++;
    CK(cudaGetLastError());
    return SPHB_OK;
}

SweepArgs sweep_args(sphb_handle *H) {
    SweepArgs A;
    A.F.x = H->x; A.F.idx = H->idx; A.F.m = H->m;
    A.F.rec = H->rec; A.F.gsum = H->gsum;
    A.a = H->a; A.du = H->du; A.dirty = H->dirty; A.wl[0] = H->wl0; A.wl[1] = H->wl1; A.wl[2] = H->wl2; A.ctl = H->d_ctl;
    A.pad = H->pad; A.na = H->na; A.m0 = H->m0; A.half_dt = H->cfg.t_i / 2;
    A.epoch = H->epoch; A.kick = 1; A.tol = H->opt.sweep_rel_tol;
    A.vn = H->vn; A.un = H->un;
    return A;
}

// SPHSimulation::step_forward (sph_simulation.cpp:55-108) in phases; between them an x-slab run
// exchanges halos with its neighbours (sphb_slab_*), a single-GPU run just goes on.
//
// The x-order check is optimistic: the kick-drift kernel raises a flag when two neighbours swapped
// places, everything after it runs regardless and writes only into the next-state buffers, and
// the flag is looked at together with the error word at the ONE host synchronisation at the end
// of the step.  In the rare step where it is set, the alive block is radix-sorted and the phases
// after the drift run again; nothing of the previous state has been overwritten by then.
int step_begin(sphb_handle *H) {   // kick-drift + order flag, wall ghosts
    H->resorts_this_step = 0;
    H->sweep_used = false;
    if (H->epoch > 0x7fff0000 - 4 * H->opt.max_sweeps) {   // stamp wrap-around
        CK(cudaMemsetAsync(H->dirty, 0, sizeof(int) * (size_t)nblk(H->cap_alive, kTile), H->stream));
        H->epoch = 8;
    }
    phase_begin(H);
    TRY(launch_begin(H, 1));
    // per-tile h maxima of a slab are gathered with atomicMax (its own tiles by the density pass, the halo's by the
    // RECORDS unpack): start the step from zero
    if (H->slab) CK(cudaMemsetAsync(H->thm, 0, sizeof(unsigned) * ((size_t)nblk(H->total, kTile) + 2), H->stream));
    TRY(launch_kick(H, 1));
    phase_end(H, 0);
    if (H->opt.sort_every_step) {
        TRY(sort_alive(H, false));
        TRY(launch_begin(H, 1));
        TRY(launch_kick(H, 0));   // ghost candidates refer to sorted slots: select again
        phase_end(H, 1);
    }
    TRY(launch_ghost_build(H, 1, 0, H->vh, H->uh, H->vh, H->uh, true));
    H->ghost_in_half = true;
    phase_end(H, 2);
    return SPHB_OK;
}

// After an attempt found the order flag set: restore the x order, re-select the ghost candidates, rebuild
// the ghosts.  The first repair of a step is a few odd-even transposition steps (a drift seldom does more
// than swap a handful of neighbours); if the next attempt still finds the flag set, the radix sort runs.
int step_resort(sphb_handle *H) {
    CK(cudaMemcpyAsync(&H->d_ctl->fallbacks, &H->fallbacks_committed, sizeof(unsigned long long),
                       cudaMemcpyHostToDevice, H->stream));
    if (H->resorts_this_step++ == 0 && H->na > 1) {
        SwapArgs W;
        W.x = H->x; W.vh = H->vh; W.uh = H->uh; W.a = H->a; W.du = H->du; W.h = H->h; W.rho = H->rho; W.P = H->P;
        W.m = H->m; W.idx = H->idx; W.pad = H->pad; W.na = H->na;
        const int grid = nblk((H->na + 1) / 2, 256);
        for (int pass = 0; pass < 2 * kSwapPasses; pass++) {
            k_swap_inversions<<<grid, 256, 0, H->stream>>>(W, pass & 1);
            H->launches++;
        }
        CK(cudaGetLastError());
        H->swap_repairs++;
    } else {
        TRY(sort_alive(H, false));
    }
    TRY(launch_begin(H, 1));
    TRY(launch_kick(H, 0));
    TRY(launch_ghost_build(H, 1, 0, H->vh, H->uh, H->vh, H->uh, true));
    H->epoch += H->opt.max_sweeps + 2;    // fresh stamps for the second attempt
    return SPHB_OK;
}

int step_density(sphb_handle *H) {
    TRY(launch_density(H, 1, 0, 1, H->Pn));
    phase_end(H, 3);
    return SPHB_OK;
}

SweepArgs step_sweep_args(sphb_handle *H) {
    SweepArgs A = sweep_args(H);
    A.a = H->an; A.du = H->dun;       // the previous a, du stay intact until step_finish
    A.Pstale = H->P; A.Pout = H->Pn;
    A.vh = H->vh; A.uh = H->uh;
    A.F.vA = H->vh; A.F.uA = H->uh;
    A.F.vB = H->vh; A.F.uB = H->uh;
    return A;
}

int launch_step_first(sphb_handle *H) {   // the first force sweep of a step, on H->stream, tiles per H->part
    SweepArgs A = step_sweep_args(H);
    if (H->opt.order == SPHB_ORDER_JACOBI) return launch_first_sel<SEL_ALIVE>(H, A);
    return launch_first_sel<SEL_INDEX>(H, A);
}

int step_force_first(sphb_handle *H) {
    TRY(sweep_plan(H));
    if (H->sweep_used) TRY(launch_sweep(H));           // the ordered sweep in one pass
    else TRY(launch_step_first(H));                    // first guess for everybody, work lists for the refinement
    phase_end(H, 4);
    return SPHB_OK;
}

int step_refine(sphb_handle *H, bool after_sweep = true) {   // runs until this handle's work list is empty
    if (H->opt.order == SPHB_ORDER_JACOBI) return SPHB_OK;
    if (H->sweep_verify_pending && after_sweep) { TRY(launch_sweep_verify(H)); H->sweep_verify_pending = false; }
    SweepArgs A = step_sweep_args(H);
    A.F.vA = H->vn; A.F.uA = H->un;   // lower-index neighbours are seen fully kicked
    TRY(launch_refine(H, A));
    phase_end(H, 5);
    return SPHB_OK;
}

// device flags of the attempt in flight: one small D2H copy + stream synchronisation
int step_flags(sphb_handle *H, int *err, int *unsorted, int *halo_changed) {
    CK(cudaMemcpyAsync(H->h_ctl, H->d_ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, H->stream));
    CK(cudaStreamSynchronize(H->stream));
    *err = H->h_ctl->err_code;
    *unsorted = H->h_ctl->unsorted;
    *halo_changed = H->h_ctl->halo_changed;
    return SPHB_OK;
}

int step_error(sphb_handle *H) {    // turn the error word just read into a status
    char b[256];
    const int code = decode_error(*H->h_ctl, H->idx, b, sizeof b);
    H->err = b;
    CK(cudaMemsetAsync(&H->d_ctl->err_code, 0, kErrBytes, H->stream));
    return code;
}

// Would work lists have (or did they have) much to do in the step just taken?  Every chunk with a non-zero kick
// would have put at least one tile on a work list.
bool sweep_busy(const sphb_handle *H) {
    const int ntiles = nblk(H->na, kTile);
    if (H->sweep_used) return H->h_ctl->sweep_moved_chunks >= ntiles / 16;
    return H->sweep_on || H->h_ctl->tiles_swept > ntiles / 8 + 16;
}

int step_finish(sphb_handle *H) {
    H->epoch += H->opt.max_sweeps + 2;
    std::swap(H->v, H->vn);
    std::swap(H->u, H->un);
    std::swap(H->h, H->hn);
    std::swap(H->rho, H->rhon);
    std::swap(H->P, H->Pn);
    std::swap(H->a, H->an);
    std::swap(H->du, H->dun);
    H->steps++;
    H->fallbacks_committed = H->h_ctl->fallbacks;
    H->refine_peak = H->h_ctl->nwork_peak;
    const int used = (H->opt.order == SPHB_ORDER_JACOBI) ? 1 : H->h_ctl->sweeps_used;
    H->sweeps_last = used;
    H->sweeps_total += used;
    // Strategy of the next step (auto mode).  The serial sweep costs the same whatever the state; the work lists cost
    // next to nothing while almost no particle feels a non-zero kick (the undisturbed lattice with dyadic
    // positions) and many times a sweep once every particle does.  The ranks of an x-slab run must all take the
    // same strategy (it decides how many messages a step exchanges): they use the agreed flag.
    if (H->opt.order == SPHB_ORDER_REFERENCE) {
        if (H->sweep_used) {
            H->sweep_steps++;
            H->sweep_flagged_last = H->h_ctl->sweep_flagged;
            H->sweep_verify_fail_last = H->h_ctl->sweep_verify_fail;
            H->sweep_burn_in_last = H->sw_B;
            // a burn-in that failed its check somewhere was too short for this state: lengthen it
            if (H->h_ctl->sweep_verify_fail > 0 && H->opt.sweep_burn_in <= 0)
                H->sweep_burn_in_last = std::min(2 * H->sw_B, 512);
        }
        H->sweep_on = H->slab ? H->agreed_busy : sweep_busy(H);
    }
    H->cfg.n_ghost = wall_ghosts(H);
    H->cfg.n_part = H->na + H->cfg.n_ghost;
    phase_collect(H);
    return SPHB_OK;
}

int step_once(sphb_handle *H) {
    TRY(step_begin(H));
    for (int attempt = 0;; attempt++) {
        TRY(step_density(H));
        TRY(step_force_first(H));
        TRY(step_refine(H));
        int err, unsorted, changed;
        TRY(step_flags(H, &err, &unsorted, &changed));
        if (unsorted && attempt < 2) {
            // whatever the kernels complained about came from the broken order: forget it, repair, redo
            CK(cudaMemsetAsync(&H->d_ctl->err_code, 0, kErrBytes, H->stream));
            TRY(step_resort(H));
            continue;
        }
        if (unsorted) { H->err = err_text(SPHB_ERR_UNSORTED); return SPHB_ERR_UNSORTED; }
        if (err) return step_error(H);
        break;
    }
    return step_finish(H);
}

int ensure_aos(sphb_handle *H, size_t n) {
    if (H->aos_cap >= n) return SPHB_OK;
    if (H->d_aos) CK(cudaFree(H->d_aos));
    H->d_aos = nullptr;
    size_t want = (size_t)H->cap_alive + 2 * (size_t)H->pad;
    if (want < n) want = n;
    CK(cudaMalloc((void **)&H->d_aos, sizeof(AosRec) * want));
    H->aos_cap = want;
    return SPHB_OK;
}

PackArgs pack_args(sphb_handle *H) {
    PackArgs A;
    A.x = H->x; A.v = H->v; A.u = H->u; A.a = H->a; A.du = H->du; A.h = H->h; A.rho = H->rho; A.P = H->P; A.m = H->m;
    A.idx = H->idx; A.slot_id = H->slot_id;
    A.gx = H->gx; A.gv = H->gv; A.gu = H->gu; A.ga = H->ga; A.gdu = H->gdu; A.gh = H->gh; A.grho = H->grho;
    A.gP = H->gP; A.gm = H->gm; A.gkey = H->gkey;
    A.vg = H->ghost_in_half ? H->vh : H->v;
    A.ug = H->ghost_in_half ? H->uh : H->u;
    A.gorder = H->gorder; A.cnt = H->cnt; A.aos = H->d_aos;
    A.out_newton = nullptr; A.out_nbr = nullptr;
    A.ctl = H->d_ctl; A.pad = H->pad; A.na = H->na; A.ng = 0;
    A.idx_base = (unsigned)H->idx_base;
    A.by_slot = H->slab ? 1 : 0;
    A.m_first = 0.0; A.ctl_rw = H->d_ctl;
    return A;
}

// make sure the ghosts' v,u are in the pads of the CURRENT v,u arrays
int ghosts_to_current(sphb_handle *H) {
    if (!H->ghost_in_half) return SPHB_OK;
    k_copy_pads<<<4, 256, 0, H->stream>>>(H->vh, H->uh, H->v, H->u, H->d_ctl, H->pad, H->na);
    H->launches++;
    CK(cudaGetLastError());
    H->ghost_in_half = false;
    return SPHB_OK;
}

int pack_to_staging(sphb_handle *H, bool with_counts, int *n_total) {
    TRY(read_ctl(H));
    const int ng = wall_ghosts(H);
    const int n = H->na + ng;
    TRY(ensure_aos(H, (size_t)n));
    PackArgs A = pack_args(H);
    A.ng = ng;
    if (with_counts) {
        if (!H->d_newton) {
            TRY(dalloc(H, &H->d_newton, (size_t)H->cap_alive + 2 * H->pad));
            TRY(dalloc(H, &H->d_nbr, (size_t)H->cap_alive + 2 * H->pad));
        }
        A.out_newton = H->d_newton; A.out_nbr = H->d_nbr;
    }
    if (n > 0) {
        k_pack<<<nblk(n, 256), 256, 0, H->stream>>>(A);
        H->launches++;
        CK(cudaGetLastError());
    }
    *n_total = n;
    return SPHB_OK;
}

}  // namespace

namespace {
SlabArgs slab_args(sphb_handle *H, int side) {
    SlabArgs A;
    A.x = H->x; A.m = H->m; A.vh = H->vh; A.uh = H->uh; A.vn = H->vn; A.un = H->un;
    A.h = H->h; A.rho = H->rho; A.P = H->P; A.idx = H->idx; A.rec = H->rec;
    A.gsum = H->gsum; A.thm = H->thm;
    A.dirty = H->dirty; A.wl[0] = H->wl0; A.wl[1] = H->wl1; A.wl[2] = H->wl2; A.ctl = H->d_ctl;
    A.pad = H->pad; A.na = H->na; A.K = H->K; A.side = side;
    return A;
}
int slab_flags(sphb_handle *H) {
    k_slab_flags<<<1, 1, 0, H->stream>>>(H->d_ctl, H->halo_l ? 1 : 0, H->halo_r ? 1 : 0);
    H->launches++;
    CK(cudaGetLastError());
    return SPHB_OK;
}
}  // namespace

namespace {
struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    std::string err;
    bool load() {
        if (lib) return true;
        for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
            lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (lib) break;
        }
        if (!lib) { err = std::string("cannot open libnccl: ") + dlerror(); return false; }
#define SPHB_SYM(F) F = reinterpret_cast<decltype(F)>(dlsym(lib, "nccl" #F)); if (!F) { err = "libnccl lacks nccl" #F; lib = nullptr; return false; }
        SPHB_SYM(GetUniqueId) SPHB_SYM(CommInitRank) SPHB_SYM(CommDestroy) SPHB_SYM(Send) SPHB_SYM(Recv)
        SPHB_SYM(AllReduce) SPHB_SYM(GroupStart) SPHB_SYM(GroupEnd) SPHB_SYM(GetErrorString)
#undef SPHB_SYM
        return true;
    }
};
NcclApi g_nccl;

#define NK(call)                                                                          \
    do {                                                                                  \
        ncclResult_t r_ = (call);                                                         \
        if (r_ != ncclSuccess) {                                                          \
            H->err = std::string(#call) + ": " + g_nccl.GetErrorString(r_);               \
            return SPHB_ERR_CUDA;                                                         \
        }                                                                                 \
    } while (0)

__global__ void k_flags(const Ctl *ctl, int *flags, int busy_if_not_swept, int ntiles, int swept) {
    flags[0] = ctl->err_code;
    flags[1] = ctl->unsorted;
    flags[2] = ctl->halo_changed;
    // the largest smoothing length this rank has seen, rounded up to float: positive floats order like ints
    flags[3] = __float_as_int(__double2float_ru(__longlong_as_double((long long)ctl->hmax_bits)));
    // the thinner of my two halos, rounded down to float and turned round (MAX over ranks = thinnest halo of the run)
    flags[4] = 0x7f800000 - __float_as_int(__double2float_rd(fmin(ctl->halo_thick[0], ctl->halo_thick[1])));
    // sweep_busy() of this rank (the strategy of the next step follows the MAX over ranks)
    flags[5] = swept ? (ctl->sweep_moved_chunks >= ntiles / 16) : (busy_if_not_swept || ctl->tiles_swept > ntiles / 8 + 16);
}

// pack both edges (one launch), one grouped send/recv with the neighbours, unpack both pads (one launch)
int native_exchange(sphb_handle *H, int what) {
    const int n = sphb_slab_message_doubles(H, what);
    SlabArgs A = slab_args(H, 0);
    if (what == MSG_BASE && !H->ghost_in_half) { A.vh = H->v; A.uh = H->u; }   // t = 0: the current v, u
    k_slab_pack2<<<dim3(std::max(1, std::min(nblk(H->K, 256), 32)), 2), 256, 0, H->stream>>>(
        A, what, H->halo_l ? H->msg_send[what][0] : nullptr, H->halo_r ? H->msg_send[what][1] : nullptr);
    H->launches++;
    CK(cudaGetLastError());
    NK(g_nccl.GroupStart());
    for (int k = 0; k < 2; k++) {
        if (!(k ? H->halo_r : H->halo_l)) continue;
        const int peer = H->rank + (k ? 1 : -1);
        NK(g_nccl.Send(H->msg_send[what][k], (size_t)n, ncclFloat64, peer, H->comm, H->stream));
        NK(g_nccl.Recv(H->msg_recv[what][k], (size_t)n, ncclFloat64, peer, H->comm, H->stream));
    }
    NK(g_nccl.GroupEnd());
    k_slab_unpack2<<<2, 1024, 0, H->stream>>>(A, what, H->halo_l ? H->msg_recv[what][0] : nullptr,
                                              H->halo_r ? H->msg_recv[what][1] : nullptr);
    H->launches++;
    CK(cudaGetLastError());
    return SPHB_OK;
}

// run fn with the handle's stream switched to the communication stream
template <class F>
int on_comm_stream(sphb_handle *H, F fn) {
    cudaStream_t keep = H->stream;
    H->stream = H->comm_stream;
    const int rc = fn();
    H->stream = keep;
    return rc;
}

// Everything of a step after the drift, up to the agreed flags (the step's one host synchronisation).
// The BASE and RECORDS exchanges run on a second stream beside the interior tiles of the density and
// force kernels; only the tiles that can see a pad ("edge": K + one tile of particles at each end)
// wait for them.
int native_attempt(sphb_handle *H, int *rounds) {
    // Two streams.  The communication stream does the BASE exchange, the density of the edge tiles, the RECORDS
    // exchange and the first sweep of the edge tiles; the main stream does the interior tiles.  Each side waits
    // for the other exactly where a tile reads its neighbour tile's results.
    CK(cudaEventRecord(H->ev_kick, H->stream));
    CK(cudaStreamWaitEvent(H->comm_stream, H->ev_kick, 0));
    TRY(on_comm_stream(H, [&]() -> int {
        TRY(native_exchange(H, MSG_BASE));
        H->part = 2;
        const int rc = launch_density(H, 1, 0, 1, H->Pn);          // edge tiles, halo in place
        H->part = 0;
        return rc;
    }));
    CK(cudaEventRecord(H->ev_edge, H->comm_stream));               // edge density done
    H->part = 1; TRY(launch_density(H, 1, 0, 1, H->Pn));            // interior tiles
    H->part = 0;
    CK(cudaEventRecord(H->ev_base, H->stream));                     // interior density done
    TRY(sweep_plan(H));
    if (H->sweep_used) {
        // The serial sweep walks the whole slab in one launch, the left halo included (it is the first segment's
        // burn-in), so it starts when the RECORDS exchange has delivered the halo's records.  What it computed for
        // the halo particles is compared with their owner's final values by the ONE ITERATE exchange that follows.
        TRY(on_comm_stream(H, [&]() -> int { return native_exchange(H, MSG_RECORDS); }));
        CK(cudaEventRecord(H->ev_rec, H->comm_stream));
        CK(cudaStreamWaitEvent(H->stream, H->ev_rec, 0));
        phase_end(H, 3);
        TRY(launch_sweep(H));
        phase_end(H, 4);
        TRY(step_refine(H));
        TRY(native_exchange(H, MSG_ITERATE));     // its unpack raises halo_changed if my estimate of a halo value was off
        *rounds = 1;
        return SPHB_OK;
    }
    TRY(on_comm_stream(H, [&]() -> int {
        TRY(native_exchange(H, MSG_RECORDS));
        CK(cudaStreamWaitEvent(H->comm_stream, H->ev_base, 0));     // edge tiles read interior records next to them
        H->part = 2;
        const int rc = launch_step_first(H);
        H->part = 0;
        return rc;
    }));
    CK(cudaEventRecord(H->ev_rec, H->comm_stream));                 // edge first sweep done
    CK(cudaStreamWaitEvent(H->stream, H->ev_edge, 0));              // interior tiles read edge records next to them
    phase_end(H, 3);
    H->part = 1; TRY(launch_step_first(H));
    H->part = 0;
    CK(cudaStreamWaitEvent(H->stream, H->ev_rec, 0));
    phase_end(H, 4);
    TRY(step_refine(H));
    for (int r = 0; r < H->fixed_rounds; r++) {
        TRY(native_exchange(H, MSG_ITERATE));
        TRY(step_refine(H));
    }
    TRY(native_exchange(H, MSG_ITERATE));     // check only: its unpack raises halo_changed if anything still moved
    *rounds = H->fixed_rounds + 1;
    return SPHB_OK;
}

// The support of the largest smoothing length ANYWHERE in the run must fit inside EVERY rank's halos: a wider one
// could reach across a halo without either rank seeing the pair.  Both numbers are agreed over all ranks (MAX of
// flags[3], flags[4]), so every rank reaches the same verdict and leaves the step with the same status: a rank
// returning alone would leave the others waiting in the next exchange for ever.
int check_halo_width(sphb_handle *H, int hmax_float_bits, int thick_code) {
    float hf, tf;
    memcpy(&hf, &hmax_float_bits, sizeof hf);
    const int tb = 0x7f800000 - thick_code;
    memcpy(&tf, &tb, sizeof tf);
    const double need = kRadius * (double)hf;
    const double have = (double)tf;
    if (need > have) {
        char b[256];
        snprintf(b, sizeof b, "thinnest x-slab halo of the run (%.6g wide) thinner than its widest kernel support (%.6g): raise halo_count", have, need);
        H->err = b;
        return SPHB_ERR_HALO;
    }
    return SPHB_OK;
}

int native_flags(sphb_handle *H) {   // -> H->h_flags[0..4] = MAX over ranks; H->h_ctl = my control block
    k_flags<<<1, 1, 0, H->stream>>>(H->d_ctl, H->d_flags, H->sweep_on ? 1 : 0, nblk(H->na, kTile), H->sweep_used ? 1 : 0);
    H->launches++;
    CK(cudaGetLastError());
    NK(g_nccl.AllReduce(H->d_flags, H->d_flags, 6, ncclInt32, ncclMax, H->comm, H->stream));
    CK(cudaMemcpyAsync(H->h_flags, H->d_flags, 6 * sizeof(int), cudaMemcpyDeviceToHost, H->stream));
    CK(cudaMemcpyAsync(H->h_ctl, H->d_ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, H->stream));
    CK(cudaStreamSynchronize(H->stream));
    H->agreed_busy = H->h_flags[5] != 0;
    return SPHB_OK;
}
}  // namespace

extern "C" {

const char *sphb_build_arch(void) { return "sm_100a"; }

void sphb_default_options(sphb_options *opt) {
    memset(opt, 0, sizeof *opt);
    opt->variable_h = 1;
    opt->fixed_h = 0.2;   // CONSTANT_H, define.hpp:29
    opt->order = SPHB_ORDER_REFERENCE;
    opt->max_sweeps = 64;
    opt->device = -1;
    opt->pad_capacity = 4096;
    opt->capacity = 0;
    opt->sort_every_step = 0;
}

static sphb_handle g_null_handle_for_errors;

const char *sphb_last_error(const sphb_handle *h) {
    return h ? h->err.c_str() : g_null_handle_for_errors.err.c_str();
}

int sphb_create(const sphb_config *cfg, const sphb_options *opt_in, sphb_handle **out) {
    sphb_handle *H = &g_null_handle_for_errors;
    if (!cfg || !out) { H->err = "null argument"; return SPHB_ERR_INVALID; }
    *out = nullptr;
    sphb_options opt;
    if (opt_in) opt = *opt_in; else sphb_default_options(&opt);
    if (cfg->n_part < 1 || cfg->n_ghost < 0 || cfg->n_ghost > cfg->n_part) { H->err = "bad n_part / n_ghost"; return SPHB_ERR_INVALID; }
    if (opt.pad_capacity < 1 || opt.max_sweeps < 1 || !(opt.sweep_rel_tol >= 0.0) || opt.sweep_mode < 0 || opt.sweep_mode > 2 ||
        (opt.sweep_group != 0 && opt.sweep_group != 8 && opt.sweep_group != 16) || opt.sweep_segment < 0 || opt.sweep_burn_in < 0) {
        H->err = "bad options";
        return SPHB_ERR_INVALID;
    }
    if (cfg->pressure_calc != SPHB_ISOTHERMAL && cfg->pressure_calc != SPHB_ADIABATIC) {
        H->err = "Unknown pressure calculation mode!";   // calculators.cpp:80-82
        return SPHB_ERR_INVALID;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        H->err = "no CUDA device: libsphb has no CPU fallback";
        return SPHB_ERR_NO_DEVICE;
    }
    H = new sphb_handle();
    H->cfg = *cfg;
    H->opt = opt;
    int dev = opt.device;
    if (dev < 0) { if (cudaGetDevice(&dev) != cudaSuccess) dev = 0; }
    H->device = dev;
    auto fail = [&](int rc) { g_null_handle_for_errors.err = H->err; sphb_destroy(H); return rc; };
#define CKC(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { H->err = std::string(#call) + ": " + cudaGetErrorString(e_); return fail(SPHB_ERR_CUDA); } } while (0)
    CKC(cudaSetDevice(dev));
    cudaDeviceProp prop;
    CKC(cudaGetDeviceProperties(&prop, dev));
    H->num_sms = prop.multiProcessorCount;
    if (prop.major < 10) {
        H->err = "libsphb is built for sm_100a only (found sm_" + std::to_string(prop.major) + std::to_string(prop.minor) + ")";
        return fail(SPHB_ERR_NO_DEVICE);
    }
    CKC(cudaStreamCreateWithFlags(&H->own_stream, cudaStreamNonBlocking));
    H->stream = H->own_stream;
    const int n_alive = cfg->n_part - cfg->n_ghost;
    H->cap_alive = std::max(opt.capacity, n_alive);
    H->pad = opt.pad_capacity;
    H->cand_cap = 2 * H->pad;
    H->total = H->cap_alive + 2 * H->pad;
    H->na = n_alive;
    H->m0 = cfg->mass;
    const size_t T = (size_t)H->total;
    double **darr[] = {&H->x, &H->v, &H->u, &H->a, &H->du, &H->h, &H->rho, &H->P, &H->m, &H->xalt, &H->vh,
                       &H->uh, &H->vn, &H->un, &H->hn, &H->rhon, &H->omn, &H->oms, &H->Pn, &H->tmp, &H->an, &H->dun};
    for (double **p : darr) { int rc = dalloc(H, p, T); if (rc) return fail(rc); }
    double **garr[] = {&H->gx, &H->gv, &H->gu, &H->ga, &H->gdu, &H->gh, &H->grho, &H->gP, &H->gm};
    for (double **p : garr) { int rc = dalloc(H, p, (size_t)H->cand_cap); if (rc) return fail(rc); }
    int rc = 0;
    rc |= dalloc(H, &H->idx, T); rc |= dalloc(H, &H->idx_alt, T); rc |= dalloc(H, &H->slot_id, T);
    rc |= dalloc(H, &H->cnt, T); rc |= dalloc(H, &H->dirty, (size_t)nblk(H->cap_alive, kTile) + 1);
    rc |= dalloc(H, &H->cand, (size_t)H->cand_cap); rc |= dalloc(H, &H->gorder, (size_t)H->cand_cap);
    rc |= dalloc(H, &H->perm, (size_t)H->cap_alive); rc |= dalloc(H, &H->iota, (size_t)H->cap_alive);
    rc |= dalloc(H, &H->gkey, (size_t)H->cand_cap); rc |= dalloc(H, &H->d_ctl, 1);
    rc |= dalloc(H, &H->wl0, (size_t)nblk(H->cap_alive, kTile) + 1); rc |= dalloc(H, &H->wl1, (size_t)nblk(H->cap_alive, kTile) + 1);
    rc |= dalloc(H, &H->wl2, (size_t)nblk(H->cap_alive, kTile) + 1);
    rc |= dalloc(H, &H->rec, T); rc |= dalloc(H, &H->gsum, T);
    rc |= dalloc(H, &H->thm, (size_t)nblk(H->total, kTile) + 2);
    H->scr_cap = ((size_t)H->cap_alive / 8 + 2) * 2 * 15;     // segments are never shorter than 8 particles
    rc |= dalloc(H, &H->scr, H->scr_cap);
    if (rc) return fail(SPHB_ERR_CUDA);
    CKC(cudaMallocHost((void **)&H->h_ctl, sizeof(Ctl)));
    memset(H->h_ctl, 0, sizeof(Ctl));
    size_t bytes = 0;
    CKC(cub::DeviceRadixSort::SortPairs(nullptr, bytes, (double *)nullptr, (double *)nullptr, (int *)nullptr,
                                        (int *)nullptr, H->cap_alive, 0, 64, H->stream));
    H->cub_bytes = bytes;
    CKC(cudaMalloc(&H->cub_tmp, std::max<size_t>(bytes, 16)));
    for (int i = 0; i <= SPHB_PHASES; i++) CKC(cudaEventCreate(&H->ev[i]));
    CKC(cudaStreamSynchronize(H->stream));
#undef CKC
    *out = H;
    return SPHB_OK;
}

void sphb_destroy(sphb_handle *H) {
    if (!H || H == &g_null_handle_for_errors) return;
    cudaSetDevice(H->device);
    if (H->stream) cudaStreamSynchronize(H->stream);
    void *ptrs[] = {H->an, H->dun, H->x, H->v, H->u, H->a, H->du, H->h, H->rho, H->P, H->m, H->xalt, H->vh, H->uh, H->vn, H->un,
                    H->hn, H->rhon, H->omn, H->oms, H->Pn, H->tmp, H->gx, H->gv, H->gu, H->ga, H->gdu, H->gh,
                    H->grho, H->gP, H->gm, H->idx, H->idx_alt, H->slot_id, H->cnt, H->dirty, H->cand, H->gorder,
                    H->perm, H->iota, H->gkey, H->d_ctl, H->cub_tmp, H->d_aos, H->d_newton, H->d_nbr,
                    H->wl0, H->wl1, H->wl2, H->rec, H->gsum, H->thm, H->scr};
    for (void *p : ptrs) if (p) cudaFree(p);
    if (H->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(H->comm);
    for (int w = 0; w < 3; w++) for (int k = 0; k < 2; k++) { if (H->msg_send[w][k]) cudaFree(H->msg_send[w][k]); if (H->msg_recv[w][k]) cudaFree(H->msg_recv[w][k]); }
    if (H->comm_stream) cudaStreamDestroy(H->comm_stream);
    for (cudaEvent_t e : {H->ev_kick, H->ev_base, H->ev_edge, H->ev_rec}) if (e) cudaEventDestroy(e);
    if (H->d_flags) cudaFree(H->d_flags);
    if (H->h_flags) cudaFreeHost(H->h_flags);
    if (H->h_ctl) cudaFreeHost(H->h_ctl);
    for (int i = 0; i <= SPHB_PHASES; i++) if (H->ev[i]) cudaEventDestroy(H->ev[i]);
    if (H->own_stream) cudaStreamDestroy(H->own_stream);
    delete H;
}

int sphb_set_stream(sphb_handle *H, void *cuda_stream) {
    if (!H) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    CK(cudaStreamSynchronize(H->stream));
    H->stream = cuda_stream ? (cudaStream_t)cuda_stream : H->own_stream;
    return SPHB_OK;
}

int sphb_get_config(const sphb_handle *H, sphb_config *out) {
    if (!H || !out) return SPHB_ERR_INVALID;
    *out = H->cfg;
    return SPHB_OK;
}

int sphb_upload(sphb_handle *H, const sphb_particle *aos, int n_part, int n_ghost) {
    if (!H) return SPHB_ERR_INVALID;
    if (!aos || n_part < 1 || n_ghost < 0 || n_ghost > n_part) { H->err = "bad upload arguments"; return SPHB_ERR_INVALID; }
    const int na = n_part - n_ghost;
    if (na > H->cap_alive || n_ghost > H->cand_cap) { H->err = "upload exceeds the handle's capacity"; return SPHB_ERR_CAPACITY; }
    CK(cudaSetDevice(H->device));
    TRY(ensure_aos(H, (size_t)n_part));
    CK(cudaMemcpyAsync(H->d_aos, aos, sizeof(AosRec) * (size_t)n_part, cudaMemcpyHostToDevice, H->stream));
    // uniform masses let every kernel take m from a register instead of a per-neighbour load; whether
    // they are is decided on the device while unpacking (a host loop over the records costs more than
    // the transfer itself at 64 Mi particles)
    const double m0 = aos[0].mass;
    H->m0 = m0;
    CK(cudaMemsetAsync(&H->d_ctl->mass_differs, 0, sizeof(int), H->stream));
    H->na = na;
    H->cfg.n_part = n_part;
    H->cfg.n_ghost = n_ghost;
    PackArgs A = pack_args(H);
    A.ng = n_ghost;
    A.m_first = m0; A.ctl_rw = H->d_ctl;
    k_unpack<<<nblk(n_part, 256), 256, 0, H->stream>>>(A);
    H->launches++;
    CK(cudaGetLastError());
    TRY(launch_begin(H, 1));
    k_check_sorted<<<nblk(na, kTile), kTile, 0, H->stream>>>(H->x, H->d_ctl, H->pad, na);
    H->launches++;
    CK(cudaGetLastError());
    TRY(read_ctl(H));
    H->uniform_mass = H->h_ctl->mass_differs == 0;
    if (H->h_ctl->unsorted) TRY(sort_alive(H, true));
    TRY(launch_ghost_build(H, 0, n_ghost, H->v, H->u, H->v, H->u));
    H->ghost_in_half = false;
    H->have_counts = false;
    return read_ctl(H);
}

int sphb_download(sphb_handle *H, sphb_particle *aos, int cap, int *n_part, int *n_ghost, int write_ids) {
    if (!H || !aos) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    int n = 0;
    TRY(pack_to_staging(H, false, &n));
    if (n > cap) { H->err = "download buffer too small"; return SPHB_ERR_CAPACITY; }
    if (write_ids) {
        CK(cudaMemcpyAsync(aos, H->d_aos, sizeof(AosRec) * (size_t)n, cudaMemcpyDeviceToHost, H->stream));
    } else {
        // leave the 8 leading bytes (const int id + padding) of every record alone
        CK(cudaMemcpy2DAsync((char *)aos + 8, sizeof(AosRec), (const char *)H->d_aos + 8, sizeof(AosRec), 80,
                             (size_t)n, cudaMemcpyDeviceToHost, H->stream));
    }
    CK(cudaStreamSynchronize(H->stream));
    const int ng = n - H->na;
    H->cfg.n_part = n;
    H->cfg.n_ghost = ng;
    if (n_part) *n_part = n;
    if (n_ghost) *n_ghost = ng;
    return SPHB_OK;
}

int sphb_download_strided(sphb_handle *H, sphb_particle *aos, int cap, int stride, int *n_alive_out, int *n_ghost_out,
                          int write_ids) {
    if (!H || !aos || stride < 1) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    int n = 0;
    TRY(pack_to_staging(H, false, &n));              // records in array-index order in the device staging buffer
    const int ng = n - H->na;
    const int na_out = (H->na + stride - 1) / stride;
    if (na_out + ng > cap) { H->err = "download buffer too small"; return SPHB_ERR_CAPACITY; }
    const size_t skip = write_ids ? 0 : 8;           // write_ids = 0: the 8 leading bytes (const int id + padding) stay
    if (na_out > 0)
        CK(cudaMemcpy2DAsync((char *)aos + skip, sizeof(AosRec), (const char *)H->d_aos + skip, sizeof(AosRec) * (size_t)stride,
                             sizeof(AosRec) - skip, (size_t)na_out, cudaMemcpyDeviceToHost, H->stream));
    if (ng > 0)
        CK(cudaMemcpy2DAsync((char *)(aos + na_out) + skip, sizeof(AosRec), (const char *)(H->d_aos + H->na) + skip, sizeof(AosRec),
                             sizeof(AosRec) - skip, (size_t)ng, cudaMemcpyDeviceToHost, H->stream));
    CK(cudaStreamSynchronize(H->stream));
    if (n_alive_out) *n_alive_out = na_out;
    if (n_ghost_out) *n_ghost_out = ng;
    return SPHB_OK;
}

int sphb_download_fields(sphb_handle *H, double *fields, int *type, int cap, int *n_part, int *n_ghost) {
    if (!H || !fields) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    int n = 0;
    TRY(pack_to_staging(H, false, &n));
    if (n > cap) { H->err = "download buffer too small"; return SPHB_ERR_CAPACITY; }
    std::vector<AosRec> tmp((size_t)n);
    CK(cudaMemcpyAsync(tmp.data(), H->d_aos, sizeof(AosRec) * (size_t)n, cudaMemcpyDeviceToHost, H->stream));
    CK(cudaStreamSynchronize(H->stream));
    for (int i = 0; i < n; i++) {
        const AosRec &r = tmp[(size_t)i];
        double *f = fields + 9 * (size_t)i;
        f[0] = r.mass; f[1] = r.pos; f[2] = r.vel; f[3] = r.acc; f[4] = r.h;
        f[5] = r.du_dt; f[6] = r.u; f[7] = r.density; f[8] = r.pressure;
        if (type) type[i] = r.type;
    }
    if (n_part) *n_part = n;
    if (n_ghost) *n_ghost = n - H->na;
    return SPHB_OK;
}

int sphb_download_counts(sphb_handle *H, int *newton_iters, int *neighbours, int cap) {
    if (!H || !newton_iters || !neighbours) return SPHB_ERR_INVALID;
    if (!H->have_counts) { H->err = "no density pass has run yet"; return SPHB_ERR_INVALID; }
    CK(cudaSetDevice(H->device));
    int n = 0;
    TRY(pack_to_staging(H, true, &n));
    if (n > cap) { H->err = "download buffer too small"; return SPHB_ERR_CAPACITY; }
    CK(cudaMemcpyAsync(newton_iters, H->d_newton, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, H->stream));
    CK(cudaMemcpyAsync(neighbours, H->d_nbr, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, H->stream));
    CK(cudaStreamSynchronize(H->stream));
    return SPHB_OK;
}

int sphb_kick_drift(sphb_handle *H) {
    if (!H) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    TRY(launch_begin(H, 1));
    TRY(launch_kick(H, 1));
    // function-level semantics: the kicked values become the particle's v, u
    std::swap(H->v, H->vh);
    std::swap(H->u, H->uh);
    TRY(read_ctl(H));
    if (H->h_ctl->unsorted) TRY(sort_alive(H, true));
    H->ghost_in_half = false;
    return SPHB_OK;
}

int sphb_ghosts(sphb_handle *H, int *n_ghost) {
    if (!H) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    TRY(launch_begin(H, 1));
    TRY(launch_kick(H, 0));
    TRY(read_ctl(H));
    if (H->h_ctl->unsorted) { H->err = err_text(SPHB_ERR_UNSORTED); return SPHB_ERR_UNSORTED; }
    TRY(launch_ghost_build(H, 1, 0, H->v, H->u, H->v, H->u));
    H->ghost_in_half = false;
    TRY(read_ctl(H));
    const int ng = wall_ghosts(H);
    H->cfg.n_ghost = ng;
    H->cfg.n_part = H->na + ng;
    if (n_ghost) *n_ghost = ng;
    return SPHB_OK;
}

int sphb_density(sphb_handle *H, int include_ghosts) {
    if (!H) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    TRY(ghosts_to_current(H));
    TRY(launch_begin(H, 0));
    TRY(launch_density(H, 1, include_ghosts ? 1 : 0, 0, H->P));
    // in place, like dc(p): the particle's h and density are replaced
    std::swap(H->h, H->hn);
    std::swap(H->rho, H->rhon);
    return read_ctl(H);
}

// records (omega, 1/h^5, ...) of every particle for its stored h, density — first half of sphb_forces
int sphb_slab_records(sphb_handle *H) {
    if (!H) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    TRY(ghosts_to_current(H));
    TRY(launch_begin(H, 0));
    // omega of every particle for its stored h, density (calc_omega), nothing else changes
    return launch_density(H, 0, 0, 0, H->P);
}
// second half of sphb_forces: the order-independent acceleration / du_dt pass, no kick
int sphb_slab_forces(sphb_handle *H) {
    if (!H) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    SweepArgs A = sweep_args(H);
    A.kick = 0;
    A.Pstale = H->P; A.Pout = H->P;
    A.vh = H->v; A.uh = H->u;
    A.F.vA = H->v; A.F.uA = H->u;
    A.F.vB = H->v; A.F.uB = H->u;
    TRY(launch_first_sel<SEL_ALL_A>(H, A));
    return read_ctl(H);
}

int sphb_forces(sphb_handle *H) {
    if (!H) return SPHB_ERR_INVALID;
    TRY(sphb_slab_records(H));
    return sphb_slab_forces(H);
}

int sphb_omega(sphb_handle *H, double *omega, int cap) {
    if (!H || !omega) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    TRY(ghosts_to_current(H));
    TRY(launch_begin(H, 0));
    TRY(launch_density(H, 0, 0, 0, H->P, true));   // omega only: h, density stay as stored
    TRY(read_ctl(H));
    const int n = H->na + wall_ghosts(H);
    if (n > cap) { H->err = "download buffer too small"; return SPHB_ERR_CAPACITY; }
    k_pack_f64<<<nblk(n, 256), 256, 0, H->stream>>>(H->omn, H->tmp, H->idx, H->gorder, H->d_ctl, H->pad, H->na,
                                                    n - H->na, (unsigned)H->idx_base, H->slab ? 1 : 0);
    H->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(omega, H->tmp, sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost, H->stream));
    CK(cudaStreamSynchronize(H->stream));
    return SPHB_OK;
}

int sphb_step(sphb_handle *H, int nsteps) {
    if (!H || nsteps < 0) return SPHB_ERR_INVALID;
    if (H->slab) { H->err = "x-slab runs step phase by phase (sphb_slab_*), with halo exchanges in between"; return SPHB_ERR_INVALID; }
    CK(cudaSetDevice(H->device));
    for (int k = 0; k < nsteps; k++) TRY(step_once(H));   // one host synchronisation per step (its flags)
    return SPHB_OK;
}

// lattice of setup.cpp:145-185 for this handle's particles (all of them, or global indices
// [idx_base, idx_base + na) of an x-slab run); no ghosts, no densities yet
int sphb_slab_init_lattice(sphb_handle *H) {
    if (!H) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    const long long n = H->slab ? H->n_global : (long long)H->na;
    if (n < 2) { H->err = "need at least 2 particles"; return SPHB_ERR_INVALID; }
    // setup.cpp:145-153,158-159, same operations in the same order
    double max_x = H->cfg.limit;
    double min_x = -max_x;
    double spacing = (max_x - min_x) / (n - 1);
    max_x -= spacing / 2;
    min_x += spacing / 2;
    const double spacing2 = (max_x - min_x) / (n - 1);
    LatticeArgs L;
    L.x = H->x; L.v = H->v; L.u = H->u; L.a = H->a; L.du = H->du; L.h = H->h; L.rho = H->rho; L.P = H->P; L.m = H->m;
    L.idx = H->idx; L.slot_id = H->slot_id; L.pad = H->pad; L.na = H->na; L.i0 = H->slab ? H->idx_base : 0;
    L.min_x = min_x; L.spacing = spacing2; L.v0 = H->cfg.v_0;
    L.u0 = (H->cfg.pressure_calc == SPHB_ADIABATIC) ? 1 / (kGamma - 1) : 0.0;
    L.h0 = H->opt.variable_h ? H->cfg.h_factor * spacing2 : H->opt.fixed_h;
    L.mass = H->cfg.mass;
    H->uniform_mass = true;
    H->m0 = H->cfg.mass;
    k_lattice<<<nblk(H->na, 256), 256, 0, H->stream>>>(L);
    H->launches++;
    CK(cudaGetLastError());
    CK(cudaMemsetAsync(H->d_ctl, 0, sizeof(Ctl), H->stream));   // no ghosts yet
    H->ghost_in_half = false;
    return slab_flags(H);
}

// v_i = +-c_s from the pre-pass density (setup.cpp:197-201)
int sphb_slab_prepass(sphb_handle *H) {
    if (!H) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    k_prepass<<<nblk(H->na, 256), 256, 0, H->stream>>>(H->x, H->v, H->u, H->rho, H->P, H->pad, H->na);
    H->launches++;
    CK(cudaGetLastError());
    return SPHB_OK;
}

// Synthetic developed start state (see k_developed): particles of this handle only, no ghosts, densities as sampled
int sphb_slab_init_developed_fill(sphb_handle *H, double t_star) {
    if (!H) return SPHB_ERR_INVALID;
    if (H->cfg.pressure_calc != SPHB_ADIABATIC || !(t_star > 0.0)) {
        H->err = "the developed start state is defined for adiabatic runs and t_star > 0";
        return SPHB_ERR_INVALID;
    }
    CK(cudaSetDevice(H->device));
    const long long n = H->slab ? H->n_global : (long long)H->na;
    if (n < 2) { H->err = "need at least 2 particles"; return SPHB_ERR_INVALID; }
    DevelopedArgs D;
    D.x = H->x; D.v = H->v; D.u = H->u; D.a = H->a; D.du = H->du; D.h = H->h; D.rho = H->rho; D.P = H->P; D.m = H->m;
    D.idx = H->idx; D.slot_id = H->slot_id; D.pad = H->pad; D.na = H->na; D.i0 = H->slab ? H->idx_base : 0; D.n_global = n;
    D.limit = H->cfg.limit; D.mass = H->cfg.mass; D.hfac = H->cfg.h_factor; D.t = t_star;
    const double L = H->cfg.limit, g = kGamma;
    D.Mhalf = 0.5 * (double)n * H->cfg.mass;
    D.rho0 = D.Mhalf / L;                              // the lattice's density
    D.P0 = (g - 1) * (1 / (g - 1)) * D.rho0;           // u0 = 1 / (gamma - 1), setup.cpp:171-172
    D.c0 = std::sqrt(g * D.P0 / D.rho0);
    D.v0 = D.c0;                                       // the adiabatic streams start at +-c_s (setup.cpp:197-201)
    D.K = D.v0 - 3.0 * D.c0;                           // v - 2 c / (gamma - 1)
    D.c_w = D.c0 - D.v0 / 3.0;
    const double rw = D.c_w / D.c0;
    D.rho_w = D.rho0 * rw * rw * rw;
    D.P_w = D.P0 * rw * rw * rw * rw * rw;
    // shock: rho0 (v0 + S) = rho1 S,  P1 = P0 + rho0 (v0 + S) v0,  e1 - e0 = (P1 + P0) / 2 (1/rho0 - 1/rho1)
    auto residual = [&](double S, double *rho1, double *P1) {
        const double r1 = D.rho0 * (D.v0 + S) / S, p1 = D.P0 + D.rho0 * (D.v0 + S) * D.v0;
        if (rho1) { *rho1 = r1; *P1 = p1; }
        return p1 / ((g - 1) * r1) - D.P0 / ((g - 1) * D.rho0) - 0.5 * (p1 + D.P0) * (1 / D.rho0 - 1 / r1);
    };
    double lo = 1e-9, hi = 100.0;
    for (int it = 0; it < 200; it++) {
        const double mid = 0.5 * (lo + hi);
        if ((residual(lo, nullptr, nullptr) > 0) == (residual(mid, nullptr, nullptr) > 0)) lo = mid; else hi = mid;
    }
    D.S = 0.5 * (lo + hi);
    residual(D.S, &D.rho1, &D.P1);
    D.x_head = (D.v0 + D.c0) * t_star;
    if (D.x_head + D.S * t_star >= L) { H->err = "t_star too late: the rarefaction has met the shock"; return SPHB_ERR_INVALID; }
    D.M1 = D.rho_w * D.c_w * t_star;
    D.M2 = D.M1 + D.rho0 * t_star * (std::pow(D.c0, 4) - std::pow(D.c_w, 4)) / std::pow(D.c0, 3);
    D.M3 = D.M2 + D.rho0 * (L - D.S * t_star - D.x_head);
    H->uniform_mass = true;
    H->m0 = H->cfg.mass;
    k_developed<<<nblk(H->na, 256), 256, 0, H->stream>>>(D);
    H->launches++;
    CK(cudaGetLastError());
    CK(cudaMemsetAsync(H->d_ctl, 0, sizeof(Ctl), H->stream));   // no ghosts yet
    H->ghost_in_half = false;
    return slab_flags(H);
}

int sphb_init_developed(sphb_handle *H, double t_star) {
    if (!H) return SPHB_ERR_INVALID;
    if (H->slab) { H->err = "x-slab runs initialise phase by phase (sphb_slab_*), with halo exchanges in between"; return SPHB_ERR_INVALID; }
    TRY(sphb_slab_init_developed_fill(H, t_star));
    TRY(sphb_ghosts(H, nullptr));
    TRY(sphb_density(H, 1));
    return sphb_forces(H);
}

int sphb_init_lattice(sphb_handle *H) {
    if (!H) return SPHB_ERR_INVALID;
    if (H->slab) { H->err = "x-slab runs initialise phase by phase (sphb_slab_*), with halo exchanges in between"; return SPHB_ERR_INVALID; }
    TRY(sphb_slab_init_lattice(H));
    if (H->cfg.pressure_calc == SPHB_ADIABATIC) {
        // pre-pass, setup.cpp:192-202
        TRY(sphb_density(H, 0));
        TRY(sphb_slab_prepass(H));
    }
    TRY(sphb_ghosts(H, nullptr));
    TRY(sphb_density(H, 1));
    return sphb_forces(H);
}

// ------------------------------------------------------------------------------------
// x-slab decomposition across the GPUs of a node.  The library does the
// device work of every phase and packs / unpacks halo messages in caller-owned device
// buffers; moving the messages between ranks (NCCL send/recv, or a device copy when the
// ranks share a process) is the caller's part — sph_b200/slabs.py.

int sphb_slab_configure(sphb_handle *H, int rank, int world, int halo_count, long long index_offset,
                        long long n_global) {
    if (!H) return SPHB_ERR_INVALID;
    if (world < 1 || rank < 0 || rank >= world || halo_count < 1 || halo_count > H->pad || index_offset < 0 ||
        n_global < H->na || index_offset + H->na > n_global || n_global >= (1ll << 31)) {
        H->err = "bad slab configuration (halo_count must fit pad_capacity; global indices must fit 31 bits)";
        return SPHB_ERR_INVALID;
    }
    CK(cudaSetDevice(H->device));
    H->slab = world > 1;
    H->rank = rank; H->world = world; H->K = halo_count;
    H->idx_base = index_offset; H->n_global = n_global;
    H->halo_l = rank > 0;
    H->halo_r = rank < world - 1;
    return slab_flags(H);
}

int sphb_slab_message_doubles(const sphb_handle *H, int what) {
    if (!H || what < 0 || what > 2 || H->K < 1) return -1;
    return msg_rows(what) * H->K + 8;
}

int sphb_slab_pack(sphb_handle *H, int what, int side, double *d_buf) {
    if (!H || !d_buf || what < 0 || what > 2 || (side != -1 && side != 1)) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    SlabArgs A = slab_args(H, side);
    if (what == MSG_BASE && !H->ghost_in_half) { A.vh = H->v; A.uh = H->u; }   // t = 0: the current v, u
    k_slab_pack<<<std::max(1, std::min(nblk(H->K, 256), 64)), 256, 0, H->stream>>>(A, what, d_buf);
    H->launches++;
    CK(cudaGetLastError());
    return SPHB_OK;
}

int sphb_slab_unpack(sphb_handle *H, int what, int side, const double *d_buf) {
    if (!H || !d_buf || what < 0 || what > 2 || (side != -1 && side != 1)) return SPHB_ERR_INVALID;
    if ((side < 0 && !H->halo_l) || (side > 0 && !H->halo_r)) { H->err = "no neighbouring slab on that side"; return SPHB_ERR_INVALID; }
    CK(cudaSetDevice(H->device));
    SlabArgs A = slab_args(H, side);
    if (what == MSG_BASE && !H->ghost_in_half) { A.vh = H->v; A.uh = H->u; }
    k_slab_unpack<<<1, 1024, 0, H->stream>>>(A, what, d_buf);
    H->launches++;
    CK(cudaGetLastError());
    return SPHB_OK;
}

int sphb_slab_step_begin(sphb_handle *H) {
    if (!H) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    return step_begin(H);
}
int sphb_slab_density(sphb_handle *H) {
    if (!H) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    return step_density(H);
}
int sphb_slab_force_first(sphb_handle *H) {
    if (!H) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    return step_force_first(H);
}
int sphb_slab_refine(sphb_handle *H) {
    if (!H) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    return step_refine(H);
}
// The flags of the attempt in flight (the one host synchronisation of a slab step):
//   flags[0] error code raised on the device (0 = none), flags[1] x-order broken by the drift,
//   flags[2] the last ITERATE unpack differed from the copy held here (sweep not known finished),
//   flags[3] the largest smoothing length seen here, as the bits of a float rounded up (orders like an int),
//   flags[4] the thinner of this rank's halos as 0x7f800000 - (bits of a float rounded down): MAX = thinnest,
//   flags[5] work lists would have much to do on this rank (the sweep strategy of the next step follows the MAX).
// Nothing is raised here: the ranks first agree (MAX over ranks) on what to do.
int sphb_slab_poll(sphb_handle *H, int *flags) {
    if (!H || !flags) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    TRY(step_flags(H, &flags[0], &flags[1], &flags[2]));
    double hm;
    { long long b = (long long)H->h_ctl->hmax_bits; memcpy(&hm, &b, sizeof hm); }
    float up = (float)hm;
    if ((double)up < hm) up = __builtin_nextafterf(up, __builtin_inff());   // round up
    memcpy(&flags[3], &up, sizeof up);
    const double th = std::min(H->h_ctl->halo_thick[0], H->h_ctl->halo_thick[1]);
    float dn = (float)th;
    if ((double)dn > th) dn = __builtin_nextafterf(dn, 0.0f);               // round down
    int tb;
    memcpy(&tb, &dn, sizeof tb);
    flags[4] = 0x7f800000 - tb;
    flags[5] = sweep_busy(H) ? 1 : 0;
    return SPHB_OK;
}
// After the ranks agreed on MAX(flags) of their polls: does the widest kernel support of the run fit inside its
// thinnest halo (the same answer on every rank), and which sweep strategy does the next step take.
int sphb_slab_agree(sphb_handle *H, const int *flags_max) {
    if (!H || !flags_max) return SPHB_ERR_INVALID;
    H->agreed_busy = flags_max[5] != 0;
    return check_halo_width(H, flags_max[3], flags_max[4]);
}
// some rank's order flag was set: sort (if mine was), forget the attempt's errors, rebuild the ghosts
int sphb_slab_resort(sphb_handle *H) {
    if (!H) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    CK(cudaMemsetAsync(&H->d_ctl->err_code, 0, kErrBytes, H->stream));
    if (H->h_ctl->unsorted) return step_resort(H);
    TRY(launch_begin(H, 1));
    TRY(launch_kick(H, 0));
    TRY(launch_ghost_build(H, 1, 0, H->vh, H->uh, H->vh, H->uh, true));
    H->epoch += H->opt.max_sweeps + 2;
    return SPHB_OK;
}
// turn the error word seen by the last poll into this handle's status and message
int sphb_slab_raise(sphb_handle *H) {
    if (!H) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    if (H->h_ctl->err_code == 0) return SPHB_OK;
    return step_error(H);
}
int sphb_slab_step_end(sphb_handle *H) {
    if (!H) return SPHB_ERR_INVALID;
    return step_finish(H);
}

// ------------------------------------------------------------------------------------
// The same protocol driven from C++: one NCCL communicator per handle, halo messages sent
// to the direct neighbours with grouped ncclSend/ncclRecv on the handle's stream, the three
// step flags agreed with one small ncclAllReduce(MAX).  A step costs the host a dozen launches
// and ONE stream synchronisation.  libnccl is opened at run time (dlopen), so libsphb.so itself
// has no link-time dependency on it.

int sphb_nccl_unique_id(void *out128) {
    sphb_handle *H = &g_null_handle_for_errors;
    if (!out128) return SPHB_ERR_INVALID;
    if (!g_nccl.load()) { H->err = g_nccl.err; return SPHB_ERR_NO_DEVICE; }
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId");
    NK(g_nccl.GetUniqueId(reinterpret_cast<ncclUniqueId *>(out128)));
    return SPHB_OK;
}

int sphb_slab_comm_init(sphb_handle *H, const void *unique_id128) {
    if (!H || !unique_id128) return SPHB_ERR_INVALID;
    if (!H->slab) { H->err = "sphb_slab_configure with world > 1 comes first"; return SPHB_ERR_INVALID; }
    if (!g_nccl.load()) { H->err = g_nccl.err; return SPHB_ERR_NO_DEVICE; }
    CK(cudaSetDevice(H->device));
    ncclUniqueId id;
    memcpy(&id, unique_id128, sizeof id);
    NK(g_nccl.CommInitRank(&H->comm, H->world, id, H->rank));
    for (int what = 0; what < 3; what++)
        for (int k = 0; k < 2; k++) {
            const size_t n = (size_t)sphb_slab_message_doubles(H, what);
            TRY(dalloc(H, &H->msg_send[what][k], n));
            TRY(dalloc(H, &H->msg_recv[what][k], n));
        }
    TRY(dalloc(H, &H->d_flags, 8));
    CK(cudaMallocHost((void **)&H->h_flags, 8 * sizeof(int)));
    // the halo exchanges and the edge tiles run beside kernels that fill the whole GPU: without a higher priority their
    // CTAs would only be scheduled when the interior kernel drains, and the exchange chain would follow it instead of hiding
    // behind it
    int prio_least = 0, prio_greatest = 0;
    CK(cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest));
    CK(cudaStreamCreateWithPriority(&H->comm_stream, cudaStreamNonBlocking, prio_greatest));
    for (cudaEvent_t *e : {&H->ev_kick, &H->ev_base, &H->ev_edge, &H->ev_rec})
        CK(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    CK(cudaStreamSynchronize(H->stream));
    return SPHB_OK;
}

// MAX over the ranks of a status code (one small all-reduce + host synchronisation): every rank returns the same status
// from a collective entry point — a rank returning alone would leave the others waiting in the next exchange
int agree_status(sphb_handle *H, int rc) {
    H->h_flags[7] = rc;
    CK(cudaMemcpyAsync(H->d_flags + 7, H->h_flags + 7, sizeof(int), cudaMemcpyHostToDevice, H->stream));
    NK(g_nccl.AllReduce(H->d_flags + 7, H->d_flags + 7, 1, ncclInt32, ncclMax, H->comm, H->stream));
    CK(cudaMemcpyAsync(H->h_flags + 7, H->d_flags + 7, sizeof(int), cudaMemcpyDeviceToHost, H->stream));
    CK(cudaStreamSynchronize(H->stream));
    const int agreed = H->h_flags[7];
    if (agreed != SPHB_OK && rc == SPHB_OK) H->err = "another slab reported an error during set-up";
    return agreed;
}

// init_particles (setup.cpp:143-228) across the slabs, halo exchanges included.  A device error on one rank does not
// stop that rank's part in the exchanges that follow (the others would wait for ever): the first status is kept,
// the phases go on, and the ranks agree on the status at the end.
int sphb_slab_init_particles(sphb_handle *H) {
    if (!H || !H->comm) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    int rc = SPHB_OK;
    auto keep = [&](int r) { if (rc == SPHB_OK && r != SPHB_OK) rc = r; };
    std::string first_err;
    auto note = [&]() { if (rc != SPHB_OK && first_err.empty()) first_err = H->err; };
    keep(sphb_slab_init_lattice(H)); note();
    if (H->cfg.pressure_calc == SPHB_ADIABATIC) {
        keep(native_exchange(H, MSG_BASE)); note();
        keep(sphb_density(H, 0)); note();
        keep(sphb_slab_prepass(H)); note();
    }
    keep(sphb_ghosts(H, nullptr)); note();
    keep(native_exchange(H, MSG_BASE)); note();
    keep(sphb_density(H, 1)); note();
    keep(sphb_slab_records(H)); note();
    keep(native_exchange(H, MSG_RECORDS)); note();
    keep(sphb_slab_forces(H)); note();
    if (rc != SPHB_OK) H->err = first_err;
    return agree_status(H, rc);
}

// the developed start state across the slabs (status agreed like sphb_slab_init_particles)
int sphb_slab_init_developed(sphb_handle *H, double t_star) {
    if (!H || !H->comm) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    int rc = SPHB_OK;
    std::string first_err;
    auto keep = [&](int r) { if (rc == SPHB_OK && r != SPHB_OK) { rc = r; first_err = H->err; } };
    keep(sphb_slab_init_developed_fill(H, t_star));
    keep(sphb_ghosts(H, nullptr));
    keep(native_exchange(H, MSG_BASE));
    keep(sphb_density(H, 1));
    keep(sphb_slab_records(H));
    keep(native_exchange(H, MSG_RECORDS));
    keep(sphb_slab_forces(H));
    if (rc != SPHB_OK) H->err = first_err;
    return agree_status(H, rc);
}

// SPHSimulation::step_forward (sph_simulation.cpp:55-108) on this rank's slab, nsteps times; every
// rank of the communicator must make the same call
int sphb_slab_step(sphb_handle *H, int nsteps) {
    if (!H || nsteps < 0 || !H->comm) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    for (int k = 0; k < nsteps; k++) {
        TRY(step_begin(H));
        int rounds = 0;
        TRY(native_attempt(H, &rounds));
        TRY(native_flags(H));
        for (int repair = 0; repair < 2 && H->h_flags[1]; repair++) {
            // some rank's drift swapped neighbours: those ranks repair their order (transposition steps first,
            // the radix sort if that was not enough), everybody repeats the attempt
            TRY(sphb_slab_resort(H));
            TRY(native_attempt(H, &rounds));
            TRY(native_flags(H));
        }
        if (H->h_flags[1]) { H->err = "x-order still broken after the sort"; return SPHB_ERR_UNSORTED; }
        while (H->h_flags[2] && !H->h_flags[0]) {
            if (rounds > H->opt.max_sweeps) { H->err = "x-slab ordered sweep did not settle"; return SPHB_ERR_NOT_CONVERGED; }
            TRY(step_refine(H));
            TRY(native_exchange(H, MSG_ITERATE));
            rounds++;
            TRY(native_flags(H));
        }
        if (H->h_flags[0]) {
            if (H->h_ctl->err_code) return step_error(H);
            H->err = "a neighbouring slab reported a device error";
            return H->h_flags[0];
        }
        TRY(check_halo_width(H, H->h_flags[3], H->h_flags[4]));
        H->rounds_last = rounds;
        TRY(step_finish(H));
    }
    return SPHB_OK;
}

int sphb_slab_rounds_last(const sphb_handle *H) { return H ? H->rounds_last : -1; }

int sphb_diagnostics(sphb_handle *H, double *out4) {
    if (!H || !out4) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    // scratch: the gather temporary is free between steps (4 * kDiagBlocks + 4 doubles are needed)
    if ((size_t)H->total < (size_t)4 * kDiagBlocks + 4) { H->err = "handle too small for the diagnostics scratch"; return SPHB_ERR_CAPACITY; }
    double *partial = H->tmp, *res = H->tmp + 4 * kDiagBlocks;
    const int nb = std::max(1, std::min(kDiagBlocks, nblk(H->na, 256)));
    k_diag_partial<<<nb, 256, 0, H->stream>>>(H->m, H->v, H->u, H->pad, H->na, partial);
    k_diag_final<<<1, 256, 0, H->stream>>>(partial, nb, res);
    H->launches += 2;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out4, res, 4 * sizeof(double), cudaMemcpyDeviceToHost, H->stream));
    CK(cudaStreamSynchronize(H->stream));
    return SPHB_OK;
}

int sphb_get_counters(sphb_handle *H, sphb_counters *out) {
    if (!H || !out) return SPHB_ERR_INVALID;
    CK(cudaSetDevice(H->device));
    memset(out, 0, sizeof *out);
    TRY(read_ctl(H));
    out->n_alive = H->na;
    out->n_ghost = wall_ghosts(H);
    out->steps = H->steps;
    out->sweeps_total = H->sweeps_total;
    out->sweeps_last = H->sweeps_last;
    out->sorts = H->sorts + H->swap_repairs;   // radix sorts + transposition repairs
    out->fallbacks = (long long)H->h_ctl->fallbacks;
    out->h_max = 0.0;
    { long long b = (long long)H->h_ctl->hmax_bits; memcpy(&out->h_max, &b, sizeof(double)); }
    out->launches = H->launches;
    unsigned long long pairs = 0, scans = 0;
    for (int b = 0; b < kStatBins; b++) { pairs += H->h_ctl->force_pairs[b]; scans += H->h_ctl->density_scans[b]; }
    out->force_pairs_mean = H->na ? (double)pairs / H->na : 0.0;
    out->density_scans_mean = (H->na + out->n_ghost) ? (double)scans / (H->na + H->h_ctl->GL + H->h_ctl->GR) : 0.0;
    out->refine_tiles = H->h_ctl->tiles_swept;
    out->refine_tile_evals = H->h_ctl->tile_evals;
    out->sweep_steps = H->sweep_steps;
    out->sweep_group = H->sweep_used ? H->sw_G : 0;
    out->sweep_segment = H->sweep_used ? H->sw_L : 0;
    out->sweep_burn_in = H->sweep_used ? H->sw_B : 0;
    out->sweep_flagged = H->sweep_used ? H->h_ctl->sweep_flagged : 0;
    out->sweep_verify_fail = H->sweep_used ? H->h_ctl->sweep_verify_fail : 0;
    if (H->have_counts && H->na > 0) {
        // reduced on the device (the per-particle counts are 8 bytes each: half a gigabyte at 64 Mi particles);
        // scratch: the gather temporary, free between steps
        unsigned long long *d_out = reinterpret_cast<unsigned long long *>(H->tmp);
        unsigned long long init[15] = {};
        init[13] = ~0ull;
        CK(cudaMemcpyAsync(d_out, init, sizeof init, cudaMemcpyHostToDevice, H->stream));
        k_count_stats<<<std::max(1, std::min(nblk(H->na, 256), 4 * H->num_sms)), 256, 0, H->stream>>>(H->cnt, H->pad, H->na, d_out);
        H->launches++;
        CK(cudaGetLastError());
        unsigned long long res[15];
        CK(cudaMemcpyAsync(res, d_out, sizeof res, cudaMemcpyDeviceToHost, H->stream));
        CK(cudaStreamSynchronize(H->stream));
        for (int k = 0; k < 12; k++) out->newton_hist[k] = (long long)res[k];
        out->nbr_mean = (double)res[12] / H->na;
        out->nbr_min = (int)res[13];
        out->nbr_max = (int)res[14];
    }
    return SPHB_OK;
}

int sphb_set_profiling(sphb_handle *H, int on) {
    if (!H) return SPHB_ERR_INVALID;
    H->profiling = on != 0;
    return SPHB_OK;
}

int sphb_get_phase_times(sphb_handle *H, double *ms, long long *launches, int reset) {
    if (!H) return SPHB_ERR_INVALID;
    for (int p = 0; p < SPHB_PHASES; p++) {
        if (ms) ms[p] = H->phase_ms[p];
        if (launches) launches[p] = H->phase_n[p];
        if (reset) { H->phase_ms[p] = 0; H->phase_n[p] = 0; }
    }
    return SPHB_OK;
}


// ------------------------------------------------------------------------------------
// sphb_density_static — the density / root-find / omega pass for callers that must not
// allocate.  It has no counterpart in the reference.
//
// Some hosts of this library cannot call cudaMalloc at the point where they need a
// density pass: applications that hand the whole device to their own arena allocator at
// start-up (a later cudaMalloc simply fails there), and code inside a latency-critical
// section where cudaMalloc's device-wide synchronisation would stall every other stream.
// Every other entry point needs a handle, and sphb_create allocates.  This one needs no
// handle: the caller passes device pointers for inputs and outputs, and the per-particle
// work arrays k_density writes besides its outputs (omega of the new and of the old
// state, packed iteration / neighbour counts) plus the control block live in ONE
// file-scope device array of fixed size, which every GPU that loads libsphb.so keeps
// until the process exits (20 MiB for the work arrays plus the control block rounded up to 256 B,
// about 1.5 KiB today; the size follows sizeof(Ctl), see kStaticCtlBytes).
//
// Rules that follow from the shared array:
//   * n is limited to SPHB_STATIC_MAX_PARTICLES; larger calls are refused, nothing runs.
//   * The array keeps what the previous call left in it.  The control block (error word,
//     fallback counter, h maximum) is cleared on the stream at the start of every call;
//     the work arrays are fully rewritten for [0, n) before anything reads them.
//   * Two calls whose device work overlaps (different streams, or different host threads)
//     race on the array.  Serialise them.
//   * The first call on each device looks up the array's device address
//     (cudaGetSymbolAddress with that device current).  Make that first call before
//     starting any stream capture.
// No cudaMalloc / cudaFree / allocating library routine is called on this path.
namespace {
// the control block's share of the array follows the struct (it holds the 64-bin statistics): the work
// arrays start at the next 256-byte boundary, so nothing the kernel writes into Ctl can reach them
constexpr size_t kStaticCtlBytes = (sizeof(Ctl) + 255) & ~(size_t)255;
static_assert(sizeof(Ctl) <= kStaticCtlBytes && kStaticCtlBytes % 256 == 0, "control block must fit its region");
constexpr size_t kStaticBytes = kStaticCtlBytes + (size_t)SPHB_STATIC_MAX_PARTICLES * (8 + 8 + 4);
// In host code this name is only a stand-in for the symbol: it is NOT a device address.
static __device__ __align__(256) unsigned char g_static_scratch[kStaticBytes];
constexpr int kMaxDevices = 64;
unsigned char *g_static_ptr[kMaxDevices] = {};   // device address of g_static_scratch, per device
std::string g_static_err;
}  // namespace

int sphb_density_static(const double *d_x, const double *d_h, const double *d_rho, int n, double mass,
                        double h_factor, double limit, int variable_h, double *d_h_out, double *d_rho_out,
                        void *cuda_stream) {
    sphb_handle *H = &g_null_handle_for_errors;   // error text is read with sphb_last_error(NULL)
    if (!d_x || !d_h || !d_rho || !d_h_out || !d_rho_out || n < 1) { H->err = "sphb_density_static: null pointer or n < 1"; return SPHB_ERR_INVALID; }
    if (n > SPHB_STATIC_MAX_PARTICLES) {
        char b[200];
        snprintf(b, sizeof b, "sphb_density_static: n = %d exceeds SPHB_STATIC_MAX_PARTICLES = %d (the static scratch "
                 "array is fixed-size; use sphb_density on a handle for larger sets)", n, SPHB_STATIC_MAX_PARTICLES);
        H->err = b;
        return SPHB_ERR_CAPACITY;
    }
    int dev = -1;
    CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) { H->err = "sphb_density_static: device ordinal out of range"; return SPHB_ERR_INVALID; }
    if (!g_static_ptr[dev]) {
        // once per device, with that device current: the symbol's address ON THIS DEVICE
        void *p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_static_scratch));
        g_static_ptr[dev] = (unsigned char *)p;
    }
    cudaStream_t st = (cudaStream_t)cuda_stream;
    unsigned char *base = g_static_ptr[dev];
    Ctl *ctl = (Ctl *)base;
    double *om_new = (double *)(base + kStaticCtlBytes);
    double *om_stale = om_new + SPHB_STATIC_MAX_PARTICLES;
    int *cnt = (int *)(om_stale + SPHB_STATIC_MAX_PARTICLES);
    CK(cudaMemsetAsync(ctl, 0, sizeof(Ctl), st));   // GL = GR = 0 (no ghosts), no error, hmax = 0
    DensityArgs A;
    A.x = d_x; A.m = nullptr; A.h = d_h; A.rho = d_rho; A.P = nullptr;
    A.hn = d_h_out; A.rhon = d_rho_out; A.omn = om_new; A.oms = om_stale; A.Pn = nullptr;
    A.rec = nullptr; A.gsum = nullptr;
    A.cnt = cnt; A.ctl = ctl; A.pad = 0; A.na = n;
    A.limit = limit; A.hfac = h_factor; A.m0 = mass;
    A.root_ghosts = 0; A.root_alive = 1; A.check_old = 0;
    A.thm = nullptr;
    const int grid = nblk(n, kTile);
    A.tile_base = 0;
    if (variable_h) k_density<true, true, false><<<grid, kTile, 0, st>>>(A);
    else k_density<false, true, false><<<grid, kTile, 0, st>>>(A);
    CK(cudaGetLastError());
    Ctl host_ctl;   // pageable stack memory: the copy allocates nothing
    CK(cudaMemcpyAsync(&host_ctl, ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (host_ctl.err_code != 0) {
        char b[256];
        const int code = decode_error(host_ctl, nullptr, b, sizeof b);   // no index array here: slot = particle
        H->err = b;
        return code;
    }
    return SPHB_OK;
}

}  // extern "C"
