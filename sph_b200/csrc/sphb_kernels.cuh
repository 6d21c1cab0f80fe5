// sphb_kernels.cuh — device code of libsphb.so (sm_100a, FP64 SIMT).
//
// Layout.  All per-particle arrays are structure-of-arrays in HBM, kept x-SORTED, with
// a pad of `pad` slots on either side of the alive block:
//
//     [ 0 .. pad-GL ) unused | [pad-GL .. pad) left ghosts | [pad .. pad+NA) alive |
//     [pad+NA .. pad+NA+GR) right ghosts | unused
//
// so every neighbour sum is a contiguous window scan over [lo, hi) = [pad-GL, pad+NA+GR)
// with no special case for ghosts (and, across GPUs, none for halo particles, which use
// the same pads).  `idx` carries the particle's ORIGINAL array index (the reference's
// sweep order, sph_simulation.cpp:94, is by array index, not by x: SURVEY.md F9); ghost
// slots have bit 31 set so they compare greater than every alive index.
//
// Arithmetic.  The file is compiled with -fmad=false: nothing is contracted implicitly,
// fused multiply-adds are written out with fma().  Where the reference's result decides
// a COUNT (ghost predicate, neighbour predicate, kick/drift that feeds them) the
// reference's expression is evaluated operation by operation with IEEE rounding.
#pragma once
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace cg = cooperative_groups;

namespace sphb {

constexpr double kEps = 1e-8;          // CALC_EPSILON, define.hpp:25
constexpr double kGamma = 5.0 / 3.0;   // GAMMA, define.hpp:32
constexpr int kMaxIterNR = 10;         // H_MAX_ITER_NR, define.hpp:39
constexpr int kMaxIterBS = 1000;       // H_MAX_ITER_BS, define.hpp:41
constexpr double kHEps = 1e-4;         // H_EPSILON, define.hpp:43
constexpr double kRadius = 2.5;        // KERNEL_RADIUS, kernel.hpp:12
constexpr double kSigma = 1.0 / 24.0;  // kernel.cpp:16
constexpr unsigned kGhostBit = 0x80000000u;

constexpr int kTile = 256;             // particles per CTA tile

enum ErrCode {
    E_OK = 0, E_CAPACITY = 3, E_NEGATIVE_H = 4, E_LOW_DENSITY = 5, E_GHOST_ZERO_H = 6,
    E_ZERO_DERIVATIVE = 7, E_NOT_FINITE = 8, E_NO_BRACKET = 9, E_NOT_CONVERGED = 10,
    E_UNSORTED = 11, E_HALO = 13
};

// Device control block (one per handle), read back by the host at sync points.
constexpr int kStatBins = 64;

struct Ctl {
    int GL, GR;                    // ghosts currently in the left / right pad
    int ncand;                     // ghost candidates appended this step
    int unsorted;                  // order check failed
    int err_code, err_unused;      // some error was raised (whichever thread came first)
    unsigned long long err_first;  // ~(((slot + 1) << 8) | code) of the LOWEST slot that raised (atomicMax; 0 = none): the
                                   // particle the reference's sequential loops would have failed at
    unsigned long long hmax_bits;  // max over every h (old and new) in [lo,hi), as ordered bits
    unsigned long long fallbacks;  // bisection fallbacks taken
    int nwork[3];                  // tiles in the work list of the sweep whose stamp is k mod 3
    int nwork_peak;                // largest work list of the current step (sizes the next step's refinement grid)
    int tile_evals;                // tile evaluations (in-tile iterations included) of the current step's refinement
    int tiles_swept;               // tiles taken from the work lists in the current step
    int sweeps_used;               // force sweeps the last step needed
    int epoch_cur;                 // stamp of the most recent force sweep (handed from launch to launch)
    int halo_left, halo_right;     // the pad on that side holds another slab's particles, not wall ghosts
    int halo_changed;              // an incoming halo iterate differed from the copy held here
    int migrated;                  // particles that changed slab in the current attempt
    double halo_thick[2];          // x distance from my edge particle to the far end of the halo on that side (+inf: wall)
    int mass_differs;              // upload: some particle's mass is not the first particle's (cleared by the upload)
    // serial sweep (k_force_sweep) of the current step
    int sweep_flagged;             // tiles it handed to the work-list refinement (reach beyond its window, index inversions)
    int sweep_verify_fail;         // segments whose burn-in did not reproduce the left segment's final values
    int sweep_moved_chunks;        // chunks holding a particle whose second kick was not exactly zero
    // statistics, one atomic per warp spread over kStatBins addresses (a single address would serialise
    // two million atomics per launch and cost more than a millisecond at 64 Mi particles)
    unsigned long long force_pairs[kStatBins];   // (i,j) pairs evaluated by the first force sweep (self included)
    unsigned long long density_scans[kStatBins]; // neighbour-window scans of the most recent density pass
};

__device__ __forceinline__ void raise(Ctl *ctl, int code, int particle) {
    atomicCAS(&ctl->err_code, 0, code);
    atomicMax(&ctl->err_first, ~((((unsigned long long)(unsigned)(particle + 1)) << 8) | (unsigned long long)code));
}

// ------------------------------------------------------------------------------------
// Arithmetic of the hot loops: the M5 quartic spline (kernel.cpp:14-64) in "r space".
// With t_k = a_k h - r (a_k = 5/2, 3/2, 1/2) and T_k = max(t_k, 0) the pieces are T_k / h, so
//     w(q)            = sigma / h^4 * [ T1^4 - 5 T2^4 + 10 T3^4 ]
//     w(q) + q w'(q)  = sigma / h^4 * [ T1^3 (a1 h - 5 r) - 5 T2^3 (a2 h - 5 r) + 10 T3^3 (a3 h - 5 r) ]
//     w'(q) / h^2     = sigma / h^5 * [ -4 T1^3 + 20 T2^3 - 40 T3^3 ]
// — no division and no q per neighbour.  An FP64 max costs three instructions on sm_100a
// (DSETP + 2 SEL); s = t + |t| = 2 max(t, 0) is ONE DADD and exact, and the powers of two it
// brings in are folded into the coefficients (exactly).
__device__ __forceinline__ double clamp2(double t) { return t + fabs(t); }   // 2 * max(t, 0)

// [T1^3 - 5 T2^3 + 10 T3^3] = -(1/4) h^5 / sigma * w'(q)/h^2 for smoothing length h at distance r >= 0.
// The FP64 pipe is what bounds the force pass, so the pieces are entered by branching on the SIGN
// of t_k (an integer test of the high word, off the FP64 pipe) instead of clamping: a neighbour in
// the outer shell costs 4 FP64 operations here, not 15.  Neighbouring lanes hold neighbouring
// particles with nearly the same h and spacing, so the branches rarely diverge.
__device__ __forceinline__ double dpoly_from(double t1, double r, double h, double k2, double k3, int hi_eps) {
#ifdef SPHB_DPOLY_BRANCHFREE
    // experiment: all three pieces unconditionally, clamped with s = t + |t| = 2 max(t, 0) (exact); straight-line code
    // that the scheduler can overlap with the reciprocal chain of the pair
    (void)hi_eps;
    const double s1 = clamp2(t1), s2 = clamp2(fma(k2, h, -r)), s3 = clamp2(fma(k3, h, -r));
    double q = (s1 * s1) * s1;
    q = fma(-5.0, (s2 * s2) * s2, q);
    q = fma(10.0, (s3 * s3) * s3, q);
    return 0.125 * q;
#endif
    double p = 0.0;
    if (__double2hiint(t1) >= hi_eps) {
        p = (t1 * t1) * t1;
        const double t2 = fma(k2, h, -r);
        if (__double2hiint(t2) >= hi_eps) {
            p = fma(-5.0, (t2 * t2) * t2, p);
            const double t3 = fma(k3, h, -r);
            if (__double2hiint(t3) >= hi_eps) p = fma(10.0, (t3 * t3) * t3, p);
        }
    }
    return p;
}
// The tests above compare the HIGH WORD of t_k with that of eps = kPieceEps * h (positive doubles order like
// their high words; a negative t_k has the sign bit set and compares below): a piece is entered only where
// t_k >= eps.  A piece with 0 <= t_k < 1e-7 h would contribute (t_k/h)^3 < 1e-21 of a typical term — far
// below the rounding of the sum — and leaving it out keeps neighbouring lanes on the same side of the
// branch when a neighbour sits on a knot of the spline to within rounding, which on a lattice with
// h = 2 spacings is EVERY neighbour (q = 0.5, 1, 1.5, 2, 2.5).
constexpr double kPieceEps = 1e-7;
constexpr double kGradC = -4.0 * kSigma;   // w'(q)/h^2 = kGradC / h^5 * dpoly(r, h)

// What the density pass leaves per particle for the force pass: everything a pair needs from
// its neighbour without a division.
struct __align__(32) Rec {
    double h;     // smoothing length
    double c5;    // -4 sigma / h^5
    double rho;   // density
    double kap;   // 1 / (rho * omega):  P / (rho^2 omega) = (gamma - 1) u kap (adiabatic), kap (isothermal)
};

struct __align__(64) Rec2 { Rec A, B; };   // new state (as seen once the sweep has passed) / old state

constexpr int kDHalo = 64;    // density pass: x staged on each side of a tile
constexpr int kFHalo = 32;    // force pass: neighbour state staged on each side of a tile
constexpr double kTinyStep = 1e-12;  // Newton step below which the sums of the previous evaluation stand for the new h
constexpr int kMergeMax = 1024;  // largest halo_count for which a cross-edge merge-split is done on the device
constexpr int kLocalIters = 48;  // cap on the in-tile iterations of one refinement sweep

// ------------------------------------------------------------------------------------
// One evaluation of the two density-pass sums at trial smoothing length h over the sorted
// window, self included:
//   Sw = sum_j [T1^4 - 5 T2^4 + 10 T3^4]           so that  sum_j w(q_j)              = sigma/h^4 Sw
//   Sg = sum_j m_j [T1^3 (a1 h - 5 r) - ...]        so that  sum_j m_j (w + q w')(q_j) = sigma/h^4 Sg
//   Sgr = sum_j m_j sign(x_i - x_j) [T1^3 - 5 T2^3 + 10 T3^3]   so that  sum_j m_j gradW_ij(h) = -4 sigma/h^5 Sgr
//   cnt = #{ j : |x_i - x_j| / h < 2.5 }            with the reference's IEEE division (kernel.cpp:31)
// sx is the staged window, sx[k] = x[wbase + k] for 0 <= k < wn, with +-INF sentinels at
// sx[-1] and sx[wn]; a scan that runs off the staged part goes on in global memory.  The scan
// stops at the first particle outside the support: the array is x-sorted, so every farther
// one is outside as well.
// one neighbour at distance r (0 <= r < 2.5 h): adds its terms to sw, sg
template <bool UM>
__device__ __forceinline__ void density_term(double r, double a1, double a2, double a3, int hi_eps, double mj,
                                             double &sw, double &sg, double &sgr) {
    const double t1 = a1 - r;
    const double c1 = (t1 * t1) * t1;
    const double r5 = 5.0 * r;
    sw = fma(c1, t1, sw);
    double g = c1 * (a1 - r5);
    double d = c1;                                 // T1^3 - 5 T2^3 + 10 T3^3: the gradient's polynomial
    const double t2 = a2 - r;
    if (__double2hiint(t2) >= hi_eps) {            // these tests run off the FP64 pipe (see dpoly_from)
        const double c2 = -5.0 * ((t2 * t2) * t2);
        sw = fma(c2, t2, sw);
        g = fma(c2, a2 - r5, g);
        d += c2;
        const double t3 = a3 - r;
        if (__double2hiint(t3) >= hi_eps) {
            const double c3 = 10.0 * ((t3 * t3) * t3);
            sw = fma(c3, t3, sw);
            g = fma(c3, a3 - r5, g);
            d += c3;
        }
    }
    if (UM) { sg += g; sgr += d; } else { sg = fma(mj, g, sg); sgr = fma(mj, d, sgr); }
}

// one side (DIR = -1 left, +1 right) of the scan around sorted slot s
template <bool UM, int DIR>
__device__ __forceinline__ void density_side(Ctl *ctl, bool own, const double *sx, int wbase, int wn, const double *__restrict__ gx,
                                             const double *__restrict__ m, int s, int lo, int hi, double xi,
                                             double h, double a1, double a2, double a3, double r_in, int hi_eps,
                                             double &sw, double &sg, double &sgr, int &c) {
    int k = s - wbase + DIR;
    double r;
#pragma unroll 1
    while (true) {                                  // staged part; the +-INF sentinels end it
        r = DIR > 0 ? (sx[k] - xi) : (xi - sx[k]);
        if (!(r < r_in)) break;
        c++;
        density_term<UM>(r, a1, a2, a3, hi_eps, UM ? 0.0 : __ldg(m + wbase + k), sw, sg, sgr);
        k += DIR;
    }
    // From here on r >= r_in (x is sorted): only particles within rounding of the support edge can
    // still count (kernel.cpp:31 decides with an IEEE division); they contribute < 1e-57 h^4.
#pragma unroll 1
    while (k >= 0 && k < wn) {
        if (!(r / h < kRadius)) return;
        c++;
        k += DIR;
        r = DIR > 0 ? (sx[k] - xi) : (xi - sx[k]);
    }
    // rare: the scan ran off the staged part (more than kDHalo neighbours on this side)
    int j = wbase + k;
#pragma unroll 1
    for (; j >= lo && j < hi; j += DIR) {
        r = DIR > 0 ? (__ldg(gx + j) - xi) : (xi - __ldg(gx + j));
        if (r < r_in) {
            c++;
            density_term<UM>(r, a1, a2, a3, hi_eps, UM ? 0.0 : __ldg(m + j), sw, sg, sgr);
        } else if (r / h < kRadius) {
            c++;
        } else {
            return;
        }
    }
    // every particle on this side lies inside the support: fine next to a wall, but on a slab
    // edge it means the halo is thinner than the kernel reaches
    if (own && (DIR < 0 ? ctl->halo_left : ctl->halo_right)) raise(ctl, E_HALO, s);
}

template <bool UM>
__device__ __forceinline__ void density_sums(Ctl *ctl, bool own, const double *sx, int wbase, int wn, const double *__restrict__ gx,
                                             const double *__restrict__ m, int s, int lo, int hi, double xi,
                                             double h, double &Sw, double &Sg, double &Sgr, int &cnt) {
    const double a1 = 2.5 * h, a2 = 1.5 * h, a3 = 0.5 * h;
    const double r_in = a1 * (1.0 - 0x1p-48);     // below: certainly inside; from there on decide exactly
    const double h2 = h * h;
    const double self = 14.375 * (h2 * h2);       // r = 0: (5/2)^4 - 5 (3/2)^4 + 10 (1/2)^4, times h^4
    double sw = self, sg = UM ? self : self * __ldg(m + s);
    int c = 1;
    double grl = 0.0, grr = 0.0;
    const int hi_eps = __double2hiint(kPieceEps * h);
    density_side<UM, -1>(ctl, own, sx, wbase, wn, gx, m, s, lo, hi, xi, h, a1, a2, a3, r_in, hi_eps, sw, sg, grl, c);
    density_side<UM, +1>(ctl, own, sx, wbase, wn, gx, m, s, lo, hi, xi, h, a1, a2, a3, r_in, hi_eps, sw, sg, grr, c);
    Sw = sw;
    Sg = sg;
    Sgr = grl - grr;                              // gradW carries sign(x_i - x_j): + for left neighbours
    cnt = c;
}

struct DensityArgs {
    const double *x, *m, *h, *rho, *P;
    double *hn, *rhon, *Pn;
    double *omn, *oms;   // omega for the new / the old (h, rho) at the new positions; may be null
    Rec2 *rec;           // force-pass records of the new and the old state; may be null
    double *gsum;        // sum_j m_j gradW_ij(h_i^new): the acceleration's own-h term; null iff rec is null
    int *cnt;            // (newton_iters << 24) | neighbour count
    Ctl *ctl;
    int pad, na;
    double limit, hfac, m0;
    int root_ghosts;     // 1: ghosts are root-found too (t = 0, setup.cpp:218-220)
    int root_alive;      // 0: omega-only pass (keeps h, rho; used by sphb_forces / sphb_omega)
    int check_old;       // 1: the old (h, rho) state is also seen by force sums (step path)
    int tile_base;       // the launch covers tiles tile_base + blockIdx.x (interior / edge launches of a slab)
    unsigned *thm;       // per tile of kTile slots: largest h (old or new) in it as float bits rounded up; may be null
};

// Fused density + variable-h root-find + omega, one thread per
// particle.  Newton follows rootfind_h (smoothing_length.cpp:191-259) and GSL's
// newton/test_delta step for step: same start value, same
// update, same stop test, same iteration cap, then the bisection fallback (:134-188).
// The last Newton evaluation is at the returned h, so its sums ARE the final density
// (calculators.cpp:45) and omega_new; the first evaluation is at the old h, so its
// derivative sum gives omega_stale = calc_omega(old h, old rho) at the new positions.
// With f(h) = m_i sigma Sw / h^5 - m_i hf / h and f'(h) = -sigma Sg / h^6 + m_i hf / h^2.
// SLAB: pads may hold another slab's particles and the launch may cover a subset of the tiles
template <bool VARH, bool UM, bool SLAB>
#ifndef SPHB_DENS_BLOCKS
#define SPHB_DENS_BLOCKS 5   // 48 registers: measured 2.43 ms at 64 M against 2.68 (6 CTAs, 40 registers, 88 B of spills) and 2.56 (4 CTAs)
#endif
__global__ void __launch_bounds__(kTile, SPHB_DENS_BLOCKS) k_density(DensityArgs A) {
    __shared__ double sxs[kTile + 2 * kDHalo + 2];
    double *sx = sxs + 1;
    Ctl *ctl = A.ctl;
    const int lo = A.pad - ctl->GL, hi = A.pad + A.na + ctl->GR;
    const int t0 = (SLAB ? (int)blockIdx.x + A.tile_base : (int)blockIdx.x) * kTile;   // tiles of [0, pad + na + pad)
    if (t0 >= hi || t0 + kTile <= lo) {
        if (A.thm && threadIdx.x == 0) A.thm[t0 / kTile] = 0u;
        return;
    }
    const int wbase = max(t0 - kDHalo, lo);
    const int wn = min(t0 + kTile + kDHalo, hi) - wbase;
    for (int k = threadIdx.x; k < wn; k += kTile) sx[k] = A.x[wbase + k];
    if (threadIdx.x == 0) { sx[-1] = -INFINITY; sx[wn] = INFINITY; }
    __syncthreads();

    const int s = t0 + threadIdx.x;
    double hloc = 0.0;
    int npass = 0;               // window scans this thread made (statistics: the algorithmic bytes of the pass)
    // pad slots that hold another slab's particles get everything from their owner
    const bool halo_slot = SLAB && ((s < A.pad && ctl->halo_left) || (s >= A.pad + A.na && ctl->halo_right));
    if (s >= lo && s < hi && !halo_slot) {
        const bool is_pad = (s < A.pad) || (s >= A.pad + A.na);
        const bool do_root = VARH && (is_pad ? A.root_ghosts : A.root_alive);
        const double xi = sx[s - wbase];
        const double mi = UM ? A.m0 : A.m[s];
        const double msig = UM ? kSigma : 1.0;   // with uniform masses Sg lacks the factor m_j = m0
        const double hold = A.h[s], rhoold = A.rho[s];
        const int npart = A.na + ctl->GL + ctl->GR;
        double hnew = hold, rhonew = rhoold, om_new = 1.0, om_stale = 1.0;
        int iters = 0, cnt = 0;
        bool tiny = false;
        bool failed = false;     // the root-find of this particle raised: the reference never reaches its density guard
        (void)tiny;
        double Sw, Sg, Sgr = 0.0;
        // 1/h, h and sum_j m_j dW/dh of the first and of the most recent evaluation: the records below are built
        // from them without further divisions by h
        double ih_last = 0.0, x_last = 0.0, dsum_last = 0.0, ih_first = 0.0, x_first = 0.0, dsum_first = 0.0;
#define SPHB_EVAL(H_, F_, DF_, DSUM_)                                                           \
    {                                                                                           \
        npass++;                                                                                \
        density_sums<UM>(ctl, SLAB && !is_pad, sx, wbase, wn, A.x, A.m, s, lo, hi, xi, (H_), Sw, Sg, Sgr, cnt); \
        const double ih_ = 1.0 / (H_), ih2_ = ih_ * ih_, ih4_ = ih2_ * ih2_;                    \
        const double rho_ = mi * (kSigma * (Sw * (ih4_ * ih_)));                                \
        F_ = rho_ - mi * A.hfac * ih_;                                                          \
        DSUM_ = UM ? -(mi * (kSigma * (Sg * (ih4_ * ih2_)))) : -(kSigma * (Sg * (ih4_ * ih2_))); \
        DF_ = DSUM_ + mi * A.hfac * ih2_;                                                       \
        rho_cur = rho_;                                                                         \
        ih_last = ih_; x_last = (H_); dsum_last = DSUM_;                                        \
    }
        (void)msig;
        if (VARH) {
            const double x0 = hold > kEps ? hold : A.hfac * (2 * A.limit / (npart - 1));
            double f, df, dsum, rho_cur;
            SPHB_EVAL(x0, f, df, dsum);
            ih_first = ih_last; x_first = x_last; dsum_first = dsum_last;
            if (A.oms) om_stale = 1.0 - dsum * (-hold / rhoold);
            if (do_root) {
                double xcur = x0;
                bool cont = true;
#pragma unroll 1
                while (cont && iters < kMaxIterNR) {
                    iters++;
                    if (df == 0.0) { raise(ctl, E_ZERO_DERIVATIVE, s); failed = true; break; }
                    const double xn = xcur - f / df;
                    if (UM && fabs(xn - xcur) <= kTinyStep * fabs(xn)) {
                        // The step did not move h beyond 1e-12 relative: this iteration passes the stop test
                        // (1e-4), and what GSL evaluates at xn is not used for another step.  The sums at xn
                        // equal the sums just taken at xcur to within 1e-12 (the density to ~1e-24 with its
                        // first-order term), five orders below the parity tolerance: do not walk the window
                        // again.  (Iteration count, returned h and stop decision are exactly the reference's.)  Uniform
                        // masses only: there dsum IS d rho_i / dh; with unequal masses (rho_i carries m_i, dsum the m_j)
                        // the window is simply walked again.
                        rho_cur = fma(-dsum, xcur - xn, rho_cur);     // rho(xn) = rho(xcur) + drho/dh (xn - xcur)
                        xcur = xn;
                        cont = false;
                        tiny = true;
                        break;
                    }
                    SPHB_EVAL(xn, f, df, dsum);
                    if (!isfinite(f) || !isfinite(df)) { raise(ctl, E_NOT_FINITE, s); failed = true; break; }
                    const double xprev = xcur;
                    xcur = xn;
                    cont = !(fabs(xcur - xprev) < kHEps * fabs(xcur) || xcur == xprev);
                }
                // (the fallback is a thousand window scans: once an error has been raised at a LOWER slot this particle's
                // cannot become the one reported, and it is skipped; a higher slot's error must not stop it — the first
                // failing particle is reported, whichever thread got there first)
                if (cont && (*((volatile unsigned long long *)&ctl->err_first) == 0ull ||
                             ((~*((volatile unsigned long long *)&ctl->err_first)) >> 8) > (unsigned long long)(s + 1))) {
                    // bisection fallback on [1e-8, 2 limit], gsl_root_test_interval rel 1e-4
                    // (its bracket end 2*limit sums over the whole domain: a slab cannot evaluate it)
                    // ... and must not try: scanning 2*limit from here would walk the whole slab, per thread
                    const bool in_slab = (ctl->halo_left | ctl->halo_right) != 0;
                    if (in_slab) { raise(ctl, E_HALO, s); failed = true; }
                    atomicAdd(&ctl->fallbacks, 1ull);
                    if (!in_slab) {
                    double xl = kEps, xh = 2 * A.limit, fl, fh, dfx, dsx;
                    int c2 = cnt;
                    SPHB_EVAL(xl, fl, dfx, dsx);
                    SPHB_EVAL(xh, fh, dfx, dsx);
                    double root = 0.5 * (xl + xh);
                    if (!isfinite(fl) || !isfinite(fh)) { raise(ctl, E_NOT_FINITE, s); failed = true; }
                    else if ((fl < 0.0 && fh < 0.0) || (fl > 0.0 && fh > 0.0)) { raise(ctl, E_NO_BRACKET, s); failed = true; }
                    else {
                        int it = 0;
                        bool go = true;
#pragma unroll 1
                        while (go && it < kMaxIterBS) {
                            it++;
                            if (fl == 0.0) { root = xl; xh = xl; }
                            else if (fh == 0.0) { root = xh; xl = xh; }
                            else {
                                const double mid = (xl + xh) / 2.0;
                                double fm;
                                SPHB_EVAL(mid, fm, dfx, dsx);
                                if (!isfinite(fm)) { raise(ctl, E_NOT_FINITE, s); failed = true; break; }
                                if (fm == 0.0) { root = mid; xl = mid; xh = mid; }
                                else if ((fl > 0.0 && fm < 0.0) || (fl < 0.0 && fm > 0.0)) {
                                    root = 0.5 * (xl + mid); xh = mid; fh = fm;
                                } else {
                                    root = 0.5 * (mid + xh); xl = mid; fl = fm;
                                }
                            }
                            double smallest = 0.0;
                            if ((xl > 0.0 && xh > 0.0) || (xl < 0.0 && xh < 0.0)) smallest = fmin(fabs(xl), fabs(xh));
                            go = !(fabs(xh - xl) < kHEps * smallest);
                        }
                    }
                    (void)c2;
                    xcur = root;
                    SPHB_EVAL(xcur, f, df, dsum);
                    }
                }
                if (xcur < 0) { raise(ctl, E_NEGATIVE_H, s); failed = true; }
                hnew = xcur;
                rhonew = rho_cur;
                if (A.omn) om_new = 1.0 - dsum * (-hnew / rhonew);
            } else {
                om_new = om_stale;   // particle keeps its stored h, rho: omega for those
            }
        } else {
            double f, df, dsum, rho_cur;
            SPHB_EVAL(hold, f, df, dsum);
            ih_first = ih_last; x_first = x_last; dsum_first = 0.0; dsum_last = 0.0;   // fixed h: omega = 1
            (void)f; (void)df; (void)dsum;
            if (is_pad ? A.root_ghosts : A.root_alive) rhonew = rho_cur;
        }
#undef SPHB_EVAL
        A.hn[s] = hnew;
        A.rhon[s] = rhonew;
        if (A.omn) { A.omn[s] = om_new; A.oms[s] = om_stale; }
        if (A.rec) {
            // c5 = -4 sigma / h^5 and kap = 1 / (rho omega) = 1 / (rho + h sum_j m_j dW/dh): the divisions the
            // force pass would otherwise make per pair.  1/h is at hand from the evaluations (one Newton step of
            // the reciprocal covers an h that moved by less than 1e-12 since its last evaluation).
            double ihn = ih_last;
            if (hnew != x_last) ihn = fma(ihn, fma(-hnew, ihn, 1.0), ihn);
            const double ihn2 = ihn * ihn;
            Rec2 r2;
            r2.A.h = hnew; r2.A.c5 = kGradC * ((ihn2 * ihn2) * ihn); r2.A.rho = rhonew;
            r2.A.kap = 1.0 / fma(hnew, (do_root ? dsum_last : dsum_first), rhonew);
            r2.B = r2.A;
            if (VARH && A.check_old) {
                const double iho = (hold == x_first) ? ih_first : 1.0 / hold, iho2 = iho * iho;
                r2.B.h = hold; r2.B.c5 = kGradC * ((iho2 * iho2) * iho); r2.B.rho = rhoold;
                r2.B.kap = 1.0 / fma(hold, dsum_first, rhoold);
            } else if (A.check_old) {
                r2.B.rho = rhoold; r2.B.kap = 1.0 / rhoold;      // fixed h: omega = 1, h unchanged
            }
            A.rec[s] = r2;
            // Sgr belongs to the LAST evaluation, which was at hnew (or at the stored h when nothing was root-found)
            A.gsum[s] = (UM ? mi : 1.0) * (r2.A.c5 * Sgr);
        }
        if (A.root_alive || A.root_ghosts) A.cnt[s] = (iters << 24) | (cnt & 0xffffff);
        if (is_pad && A.Pn != A.P) A.Pn[s] = A.P[s];
        // ensure_nonzero_density (calculators.cpp:174-186) guards every particle a force sum
        // touches, in whichever state (new, or old when check_old) it will be seen
        if (!failed && (rhonew < kEps || (A.check_old && rhoold < kEps))) raise(ctl, E_LOW_DENSITY, s);
        hloc = fmax(hnew, hold);
    }
    // window scans of this tile (one atomic per warp)
    {
        const int np = __reduce_add_sync(0xffffffffu, npass);
        if ((threadIdx.x & 31) == 0 && np)
            atomicAdd(&ctl->density_scans[(blockIdx.x * (kTile / 32) + (threadIdx.x >> 5)) % kStatBins], (unsigned long long)np);
    }
    // Block max of h -> global upper bound for the force pass window.  It is only a scan bound, so it
    // may be rounded UP to single precision: positive floats order like their bit patterns, which makes
    // the warp step one REDUX instruction instead of five 64-bit shuffle / select rounds.
    __shared__ unsigned smax[kTile / 32];
    const unsigned wmax = __reduce_max_sync(0xffffffffu, __float_as_uint(__double2float_ru(hloc)));
    if ((threadIdx.x & 31) == 0) smax[threadIdx.x >> 5] = wmax;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned mx = smax[0];
        for (int k = 1; k < kTile / 32; k++) mx = max(mx, smax[k]);
        const unsigned long long bits = (unsigned long long)__double_as_longlong((double)__uint_as_float(mx));
        // after the first CTAs the bound rarely grows: look before paying for an atomic on one hot address
        if (bits > *((volatile unsigned long long *)&ctl->hmax_bits)) atomicMax(&ctl->hmax_bits, bits);
        // the serial sweep bounds its neighbour scan by the largest h of the tiles around a particle
        // (a slab's halo slots are skipped above: their maximum arrives with the RECORDS message)
        if (A.thm) { if (SLAB) atomicMax(A.thm + t0 / kTile, mx); else A.thm[t0 / kTile] = mx; }
    }
}

// ------------------------------------------------------------------------------------
// Force pass: acceleration (calculators.cpp:63-129) and du/dt (:192-213) fused, one
// thread per alive particle, neighbours "as seen by i" (SURVEY.md Appendix A.6):
// state A (new h, rho, omega and the current iterate of the fully-kicked v, u) for
// particles the reference's sweep has already updated when it reaches i, state B
// (last step's h, rho, omega_stale, half-kicked v, u) for the others and for ghosts.
struct ForceArrays {
    const double *x;
    const unsigned *idx;
    const double *m;
    const Rec2 *rec;
    const double *gsum;        // sum_j m_j gradW_ij(h_i) from the density pass
    const double *vA, *uA;     // current iterate of the fully kicked v, u (read with .cg: other CTAs update it)
    const double *vB, *uB;     // half-kicked base state
};

enum { SEL_INDEX = 0, SEL_ALIVE = 1, SEL_ALL_A = 2 };

struct SweepArgs {
    ForceArrays F;
    const double *Pstale;   // stored pressure of the previous force pass
    double *Pout;           // new stored pressure
    double *a, *du;         // outputs
    double *vn, *un;        // fully-kicked outputs (iterate)
    const double *vh, *uh;  // half-kicked base state of i itself
    int *dirty;             // per-tile stamp: epoch of the last sweep that changed something in range
    int *wl[3];             // work lists of tiles, by sweep stamp mod 3
    Ctl *ctl;
    int pad, na;
    double m0, half_dt;
    int epoch;              // stamp of this sweep
    int kick;               // 0: t = 0 semantics (no kick, calculators only)
    double tol;             // 0: iterate to the bitwise fixed point; > 0: until every change is below tol (relative)
    int tile_base;          // first sweep: the launch covers tiles tile_base + blockIdx.x
};

// per-thread invariants of one force evaluation
template <bool ADIA>
struct Own {
    double x, h, rho, v, PrRho, c_ac, c_ec, P;
    unsigned idx;
    int hi_eps;      // high word of kPieceEps * h: spline pieces below it are left out (see dpoly_from)
    __device__ __forceinline__ void set(double xi, unsigned idxi, const Rec &r, double vi, double ui, double Pstale) {
        x = xi; idx = idxi; h = r.h; rho = r.rho; v = vi;
        hi_eps = __double2hiint(kPieceEps * r.h);
        P = ADIA ? ((kGamma - 1) * ui) * r.rho : r.rho;                 // calculators.cpp:131-137
        c_ac = ADIA ? sqrt(kGamma * Pstale / r.rho) : 1.0;              // stale stored pressure
        c_ec = ADIA ? sqrt(kGamma * P / r.rho) : 1.0;                   // fresh pressure
        PrRho = ADIA ? ((kGamma - 1) * ui) * r.kap : r.kap;             // P_i / (rho_i^2 omega_i)
    }
};

// One (i, j) pair; ar = |x_i - x_j| >= 0 and the sign of the gradient is applied by the caller per
// side.  The acceleration's first term, gradW(h_i) P_i/(rho_i^2 om_i), does not depend on j's state:
// its sum over j comes from the density pass (gsum).  accA gathers m_j [gradW(h_j) P_j/(rho_j^2 om_j)
// + gradW(hbar) Pi_ij(c_ac)], accE gathers m_j v_ij gradW(hbar) [P_i/(rho_i^2 om_i) + Pi_ij(c_ec)/2],
// both without the sign.  dir = +1 for a left neighbour (x_i - x_j = +ar), -1 for a right one.
// Everything is written with hs = h_i + h_j = 2 hbar.
//
// The pair is evaluated in two parts.  pair_geom: everything that depends on positions, smoothing
// lengths, densities and omegas only (the reciprocal, both spline polynomials) — four numbers.
// pair_apply: the part that needs the neighbour's velocity and energy, a dozen operations.  The
// ordered sweep changes only (v_j, u_j) of a neighbour while it runs, so the serial sweep kernel
// does the first part for all pairs of a chunk in parallel and only the second part along the
// dependency chain.  Every force kernel goes through these two functions, in the same order of
// accumulation (farthest neighbour first, per side), so a particle evaluated by any of them from
// the same inputs gets the same bits.
struct PairK {
    double K2;     // mu_ij = K2 (v_i - v_j):  hbar r_ij / (r^2 + 0.01 hbar^2), signed
    double irho;   // 1 / ((rho_i + rho_j) / 2)
    double gb;     // |gradW(hbar)|
    double G1;     // |gradW(h_j)| / (rho_j omega_j) (x (gamma - 1) when adiabatic): times u_j (adiabatic) it is the P_j term
};

template <bool ADIA>
__device__ __forceinline__ bool pair_geom(const Own<ADIA> &o, double ar, double dir, const Rec &rj, PairK &k) {
    const double hs = o.h + rj.h;
    const double t1j = fma(2.5, rj.h, -ar), t1b = fma(1.25, hs, -ar);
    if (max(__double2hiint(t1j), __double2hiint(t1b)) < o.hi_eps) return false;   // beyond (or on the edge of) 2.5 h_j and 2.5 hbar
    const double rs = o.rho + rj.rho;
    const double D = fma(ar, ar, 0.0025 * (hs * hs));        // r^2 + 0.01 hbar^2
    // one reciprocal serves 1/hbar, 1/D and 1/(rho_i + rho_j)
    const double DP = D * rs;
    const double R = 1.0 / (hs * DP);
    const double ihs = R * DP;
    const double t = R * hs;              // 1 / (D rs)
    const double iD = t * rs;
    const double irs = t * D;
    const double ihs2 = ihs * ihs;
    // gradW(hbar) without its sign: -4 sigma / hbar^5 * poly = -128 sigma / hs^5 * poly
    k.gb = dpoly_from(t1b, ar, hs, 0.75, 0.25, o.hi_eps) * ((32.0 * kGradC) * ((ihs2 * ihs2) * ihs));
    k.G1 = dpoly_from(t1j, ar, rj.h, 1.5, 0.5, o.hi_eps) * (rj.c5 * (ADIA ? (kGamma - 1) * rj.kap : rj.kap));
    k.K2 = ((0.5 * hs) * (dir * ar)) * iD;
    k.irho = 2.0 * irs;                   // 1 / ((rho_i + rho_j) / 2)
    return true;
}

template <bool ADIA, bool UM>
__device__ __forceinline__ void pair_apply(const Own<ADIA> &o, const PairK &k, double vj, double uj, double mj,
                                           double &accA, double &accE) {
    const double v_ij = o.v - vj;
    const double mu = k.K2 * v_ij;        // same sign as v_ij . r_ij
    // artificial viscosity (calculators.cpp:148-172): alpha = 1, beta = 2, eta^2 = 0.01 hbar^2; none for a receding pair
    double visc_ac = 0.0, visc_ec = 0.0;
    if (!(mu > 0)) {
        const double two_mu2 = 2.0 * (mu * mu);
        visc_ac = fma(-o.c_ac, mu, two_mu2) * k.irho;
        visc_ec = fma(-o.c_ec, mu, two_mu2) * k.irho;
    }
    const double gj = ADIA ? k.G1 * uj : k.G1;
    const double ta = fma(k.gb, visc_ac, gj);
    const double te = (v_ij * k.gb) * fma(0.5, visc_ec, o.PrRho);
    if (UM) { accA += ta; accE += te; }
    else { accA = fma(mj, ta, accA); accE = fma(mj, te, accE); }
}

// Shared-memory image of a tile and its surroundings for the force pass.
struct ForceSmem {
    double x[kTile + 2 * kFHalo + 2];
    // records split in 16-byte halves so that a warp's LDS.128 is conflict-free; [0] = new state A, [1] = old state B
    double2 hc[2][kTile + 2 * kFHalo];   // (h, c5)
    double2 rk[2][kTile + 2 * kFHalo];   // (rho, kap)
    double vA[kTile + 2 * kFHalo], uA[kTile + 2 * kFHalo];
    double vB[kTile + 2 * kFHalo], uB[kTile + 2 * kFHalo];
    unsigned idx[kTile + 2 * kFHalo];
    double gsum[kTile], pst[kTile];      // own particles: gradient sum of the density pass, stored pressure
    unsigned hw[kTile / 32];             // per warp: largest smoothing length it staged (float bits, rounded up)
};

template <bool VOL>
__device__ __forceinline__ double ldv(const double *p) { return VOL ? __ldcg(p) : __ldg(p); }

template <int SEL>
__device__ __forceinline__ bool sees_new_state(unsigned idxj, unsigned idxi) {
    if (SEL == SEL_INDEX) return idxj < idxi;            // the reference's sweep has already updated j
    if (SEL == SEL_ALIVE) return !(idxj & kGhostBit);
    return true;
}

// All neighbours on one side (DIR = -1: left, x_j <= x_i; +1: right) of particle `o`, summed from the
// FARTHEST one inwards (the order in which the serial sweep meets them): first the extent of the scan is
// found, then the pairs are evaluated on the way back.  gradW carries sign(x_i - x_j)
// (calculators.cpp:17-27): the unsigned sums of this side enter a with -DIR... i.e. a -= accA on the left,
// a += accA on the right, and du/dt with the opposite signs.  `jend` receives the farthest slot within reach.
// SPLIT (refinement): a_sum / e_sum receive this side's unsigned sums instead of being merged, and `dyn` tells
// whether the side holds a neighbour whose value can change while the tile iterates (seen in its new state and
// inside the tile [dyn_lo, dyn_hi) of the staged window): a side without one need not be recomputed.
template <int SEL, bool ADIA, bool UM, bool FIRST, int DIR, bool SPLIT = false>
__device__ __forceinline__ void force_side(const SweepArgs &A, const ForceSmem &S, const Own<ADIA> &o, int ks,
                                           int wbase, int wn, int lo, int hi, double reach_win,
                                           double &a_sum, double &e_sum, int &npairs, int &jend,
                                           int dyn_lo = 0, int dyn_hi = 0, bool *dyn = nullptr) {
    const double *sx = S.x + 1;
    const double dir = DIR > 0 ? -1.0 : 1.0;      // sign of x_i - x_j
    double accA = 0.0, accE = 0.0;
    // how far the scan goes inside the staged window: up to the first particle beyond reach_win (or a sentinel)
    int kend = ks + DIR;
#pragma unroll 1
    while ((DIR > 0 ? (sx[kend] - o.x) : (o.x - sx[kend])) < reach_win) kend += DIR;
    int jfar = wbase + kend - DIR;                // farthest slot within reach so far (ks itself if none)
    // Inside the staged window no particle has a larger h than reach_win / 2.5, so whatever lies beyond the
    // point where the scan stopped cannot interact.  BEYOND the window only the global bound is known: if it
    // reaches past the outermost staged particle, the scan goes on from there in global memory (rare: it takes
    // an h several times the local one, or more than kFHalo neighbours on one side) — those are the farthest
    // neighbours, so they are summed first.
    const double dist_end = DIR > 0 ? (sx[wn - 1] - o.x) : (o.x - sx[0]);
    // (the global bound is re-read here rather than kept in registers through the loops)
    const double reach = kRadius * __longlong_as_double(*((volatile long long *)&A.ctl->hmax_bits));
    if (dist_end < reach) {
        const int jstart = DIR > 0 ? wbase + wn : wbase - 1;
        int j = jstart;
#pragma unroll 1
        for (; j >= lo && j < hi; j += DIR) {
            const double xj = __ldg(A.F.x + j);
            if (!((DIR > 0 ? (xj - o.x) : (o.x - xj)) < reach)) break;
        }
        if ((j < lo || j >= hi) && (DIR < 0 ? A.ctl->halo_left : A.ctl->halo_right)) raise(A.ctl, E_HALO, wbase + ks);
        if (j != jstart) jfar = j - DIR;
#pragma unroll 1
        for (int q = j - DIR; DIR > 0 ? q >= jstart : q <= jstart; q -= DIR) {
            const double xj = __ldg(A.F.x + q);
            const double ar = DIR > 0 ? (xj - o.x) : (o.x - xj);
            const bool useA = sees_new_state<SEL>(__ldg(A.F.idx + q), o.idx);
            const Rec rj = useA ? A.F.rec[q].A : A.F.rec[q].B;
            PairK pk;
            if (pair_geom<ADIA>(o, ar, dir, rj, pk)) {
                const double vj = (useA && !FIRST) ? ldv<true>(A.F.vA + q) : __ldg(A.F.vB + q);
                const double uj = (useA && !FIRST) ? ldv<true>(A.F.uA + q) : __ldg(A.F.uB + q);
                pair_apply<ADIA, UM>(o, pk, vj, uj, UM ? 0.0 : __ldg(A.F.m + q), accA, accE);
                npairs++;
            }
        }
    }
    // the staged part, inwards
#pragma unroll 1
    for (int k = kend - DIR; k != ks; k -= DIR) {
        const double ar = DIR > 0 ? (sx[k] - o.x) : (o.x - sx[k]);
        const bool useA = sees_new_state<SEL>(S.idx[k], o.idx);
        const int sel = useA ? 0 : 1;
        const double2 hc = S.hc[sel][k], rk = S.rk[sel][k];
        Rec rj;
        rj.h = hc.x; rj.c5 = hc.y; rj.rho = rk.x; rj.kap = rk.y;
        PairK pk;
        if (pair_geom<ADIA>(o, ar, dir, rj, pk)) {
            const double vj = (useA && !FIRST) ? S.vA[k] : S.vB[k];
            const double uj = (useA && !FIRST) ? S.uA[k] : S.uB[k];
            pair_apply<ADIA, UM>(o, pk, vj, uj, UM ? 0.0 : __ldg(A.F.m + wbase + k), accA, accE);
            npairs++;
            if (SPLIT && useA && k >= dyn_lo && k < dyn_hi) *dyn = true;
        }
    }
    jend = jfar;
    if (SPLIT) {
        a_sum = accA;
        e_sum = accE;
    } else {
        a_sum = DIR > 0 ? (a_sum + accA) : (a_sum - accA);
        e_sum = DIR > 0 ? (e_sum - accE) : (e_sum + accE);
    }
}

// Every tile holding a particle within reach of a changed one must be looked at again next sweep
// (all but `skip_tile`).  The converged lanes of a warp queue together: one lane does the atomics for
// all of them (a step in which every particle moves would otherwise drown in atomics).
__device__ __forceinline__ void queue_dependents(const SweepArgs &A, bool changed, int jmin, int jmax, int epoch,
                                                 int skip_tile) {
    const unsigned act = __activemask();
    if (!__ballot_sync(act, changed)) return;
    const int wmin = __reduce_min_sync(act, changed ? jmin : 0x7fffffff);
    const int wmax = __reduce_max_sync(act, changed ? jmax : -1);
    if ((threadIdx.x & 31) != __ffs(act) - 1) return;
    const int ntiles = (A.na + kTile - 1) / kTile;
    int ta = max((wmin - A.pad) / kTile, 0);
    const int tb = min((wmax - A.pad) / kTile, ntiles - 1);
    if (wmin < A.pad) ta = 0;
    for (int t = ta; t <= tb; t++) {
        if (t == skip_tile) continue;
        if (atomicMax(A.dirty + t, epoch) < epoch) {
            const int slot = atomicAdd(&A.ctl->nwork[(epoch + 1) % 3], 1);
            A.wl[(epoch + 1) % 3][slot] = t;
        }
    }
}

// Did a kicked value move?  tol == 0: any bit; else by more than tol relative to `scale`.
__device__ __forceinline__ bool moved_by(double a, double b, double tol, double scale) {
    if (tol == 0.0) return __double_as_longlong(a) != __double_as_longlong(b);
    return fabs(a - b) > tol * scale;
}

// Evaluate one tile: stage, compute a, du/dt (and the kick), store.  Returns through
// `changed_any` whether any particle's kicked (v, u) differs from what was stored before.
// FIRST: sweep 1 of a step (vA == vB: everybody is seen half-kicked, outputs always stored).
template <int SEL, bool ADIA, bool UM, bool FIRST>
__device__ __forceinline__ void force_tile(const SweepArgs &A, int tile, int lo, int hi, double Hb,
                                           int epoch, bool &changed_any, int &npairs_out) {
    __shared__ ForceSmem S;
    const int t0 = A.pad + tile * kTile;
    const int tend = min(t0 + kTile, A.pad + A.na);
    const int wbase = max(t0 - kFHalo, lo);
    const int wn = min(t0 + kTile + kFHalo, hi) - wbase;
    double *sx = S.x + 1;
    __syncthreads();                                   // previous tile's readers are done
    double hstaged = 0.0;
    for (int k = threadIdx.x; k < wn; k += kTile) {
        const int j = wbase + k;
        sx[k] = A.F.x[j];
        S.idx[k] = A.F.idx[j];
        {
            const Rec2 r2 = A.F.rec[j];
            hstaged = fmax(hstaged, fmax(r2.A.h, r2.B.h));
            S.hc[0][k] = make_double2(r2.A.h, r2.A.c5); S.rk[0][k] = make_double2(r2.A.rho, r2.A.kap);
            S.hc[1][k] = make_double2(r2.B.h, r2.B.c5); S.rk[1][k] = make_double2(r2.B.rho, r2.B.kap);
        }
        S.vB[k] = __ldg(A.F.vB + j);
        S.uB[k] = __ldg(A.F.uB + j);
        if (!FIRST) { S.vA[k] = ldv<true>(A.F.vA + j); S.uA[k] = ldv<true>(A.F.uA + j); }
    }
    if (t0 + (int)threadIdx.x < tend) {   // issued with the staging loads, read after the pair loops
        S.gsum[threadIdx.x] = A.F.gsum[t0 + threadIdx.x];
        S.pst[threadIdx.x] = A.Pstale[t0 + threadIdx.x];
    }
    {
        const unsigned wm = __reduce_max_sync(0xffffffffu, __float_as_uint(__double2float_ru(hstaged)));
        if ((threadIdx.x & 31) == 0) S.hw[threadIdx.x >> 5] = wm;
    }
    if (threadIdx.x == 0) { sx[-1] = -INFINITY; sx[wn] = INFINITY; }
    __syncthreads();
    // the scan bound inside the staged window: 2.5 x the largest h staged (rounded up), never above the global one
    unsigned hwb = S.hw[0];
#pragma unroll
    for (int w = 1; w < kTile / 32; w++) hwb = max(hwb, S.hw[w]);
    const double reach_win = fmin(kRadius * Hb, kRadius * (double)__uint_as_float(hwb));

    const int s = t0 + threadIdx.x;
    bool changed = false;
    int npairs = 0;

    if constexpr (FIRST) {
        // sweep 1 of a step (or the kick-less t = 0 pass): everybody is seen half-kicked, one evaluation.
        // (Everything is scoped inside this branch on purpose: hoisting ks / reach out of it costs the
        // 64-register kernel 100 bytes of extra spill traffic per thread and 10 % of its speed.)
        if (s < tend) {
            const int ks2 = s - wbase;
            Own<ADIA> o;
            const double vi = S.vB[ks2], ui = S.uB[ks2];      // = A.vh[s], A.uh[s]: staged with the tile
            {
                Rec ro;
                ro.h = S.hc[0][ks2].x; ro.c5 = S.hc[0][ks2].y; ro.rho = S.rk[0][ks2].x; ro.kap = S.rk[0][ks2].y;
                o.set(sx[ks2], S.idx[ks2], ro, vi, ui, S.pst[threadIdx.x]);
            }
            double a_sum = 0.0, e_sum = 0.0;
            int jmin = s, jmax = s;
            npairs = 1;   // self: its energy term is evaluated by the reference too (and is exactly 0)
            force_side<SEL, ADIA, UM, true, -1>(A, S, o, ks2, wbase, wn, lo, hi, reach_win, a_sum, e_sum, npairs, jmin);
            force_side<SEL, ADIA, UM, true, +1>(A, S, o, ks2, wbase, wn, lo, hi, reach_win, a_sum, e_sum, npairs, jmax);
            if (UM) { a_sum *= A.m0; e_sum *= A.m0; }
            a_sum = fma(-o.PrRho, S.gsum[threadIdx.x], a_sum);   // - P_i/(rho_i^2 om_i) sum_j m_j gradW_ij(h_i)
            if (!A.kick) {
                A.a[s] = a_sum;
                A.du[s] = e_sum;
                A.Pout[s] = o.P;
            } else {
                // second half-kick, sph_simulation.cpp:105-106, operation by operation
                const double vf = __dadd_rn(vi, __dmul_rn(a_sum, A.half_dt));
                const double uf = __dadd_rn(ui, __dmul_rn(e_sum, A.half_dt));
                A.a[s] = a_sum; A.du[s] = e_sum; A.Pout[s] = o.P;
                A.vn[s] = vf; A.un[s] = uf;
                // later-index neighbours used (vh, uh) as their guess for this particle
                changed = moved_by(vf, vi, A.tol, fmax(fabs(vi), o.c_ec)) || moved_by(uf, ui, A.tol, fabs(ui));
                if (SEL == SEL_INDEX) queue_dependents(A, changed, jmin, jmax, epoch, -1);
            }
        }
    } else {
        // A refinement sweep iterates the tile ON ITS SHARED-MEMORY IMAGE until nothing in it moves any
        // more (the particles outside the tile keep their staged values): the chain of dependencies
        // inside a tile then costs CTA barriers instead of grid-wide sweeps.
        const bool active = s < tend;
        const int ks = s - wbase;
        bool local_cap = false;
        Own<ADIA> o;
        double vi = 0.0, ui = 0.0, gsum = 0.0, vscale = 1.0;
        if (active) {
            vi = S.vB[ks]; ui = S.uB[ks]; gsum = S.gsum[threadIdx.x];
            Rec ro;
            ro.h = S.hc[0][ks].x; ro.c5 = S.hc[0][ks].y; ro.rho = S.rk[0][ks].x; ro.kap = S.rk[0][ks].y;
            o.set(sx[ks], S.idx[ks], ro, vi, ui, S.pst[threadIdx.x]);
            vscale = fmax(fabs(vi), o.c_ec);   // velocities pass through zero: judge them against max(|v|, c_s)
        }
        double a_sum = 0.0, e_sum = 0.0, vf = 0.0, uf = 0.0;
        int jmin = s, jmax = s;
        // Per side: its unsigned sums and whether anything in it can move while this tile iterates.  With the
        // array order equal to the x order (the usual case) the right-hand neighbours all have a higher index,
        // are seen in their old state, and that half of the work is done once per sweep.
        double La = 0.0, Le = 0.0, Ra = 0.0, Re = 0.0;
        bool dynL = false, dynR = false;
        const int dyn_lo = t0 - wbase, dyn_hi = tend - wbase;
#pragma unroll 1
        for (int it = 0;; it++) {
            bool moved = false;
            if (active) {
                if (it == 0 || dynL)
                    force_side<SEL, ADIA, UM, false, -1, true>(A, S, o, ks, wbase, wn, lo, hi, reach_win, La, Le, npairs, jmin,
                                                               dyn_lo, dyn_hi, &dynL);
                if (it == 0 || dynR)
                    force_side<SEL, ADIA, UM, false, +1, true>(A, S, o, ks, wbase, wn, lo, hi, reach_win, Ra, Re, npairs, jmax,
                                                               dyn_lo, dyn_hi, &dynR);
                // merged exactly as the first sweep merges them: (0 - left) + right, (0 + left) - right
                a_sum = (0.0 - La) + Ra;
                e_sum = (0.0 + Le) - Re;
                if (UM) { a_sum *= A.m0; e_sum *= A.m0; }
                a_sum = fma(-o.PrRho, gsum, a_sum);
                vf = __dadd_rn(vi, __dmul_rn(a_sum, A.half_dt));
                uf = __dadd_rn(ui, __dmul_rn(e_sum, A.half_dt));
                moved = moved_by(vf, S.vA[ks], A.tol, vscale) || moved_by(uf, S.uA[ks], A.tol, fabs(ui));
            }
            if (!__syncthreads_or(moved)) break;      // also: everybody is done reading the image
            if (moved) { S.vA[ks] = vf; S.uA[ks] = uf; }
            if (threadIdx.x == 0) atomicAdd(&A.ctl->tile_evals, 1);
            __syncthreads();
            if (it >= kLocalIters) { local_cap = true; break; }
        }
        if (threadIdx.x == 0) { atomicAdd(&A.ctl->tile_evals, 1); atomicAdd(&A.ctl->tiles_swept, 1); }
        if (active) {
            const double vo = __ldcg(A.vn + s), uo = __ldcg(A.un + s);
            changed = moved_by(vf, vo, A.tol, vscale) || moved_by(uf, uo, A.tol, fabs(ui));
            // a, du feed the next step's first kick: they must come from THIS (possibly final)
            // evaluation even when the kicked values round to what is already stored
            A.a[s] = a_sum; A.du[s] = e_sum;
            if (changed) { __stcg(A.vn + s, vf); __stcg(A.un + s, uf); }
            // this tile itself only if its local iteration was cut short
            if (SEL == SEL_INDEX) queue_dependents(A, changed, jmin, jmax, epoch, local_cap ? -1 : tile);
        }
    }
    changed_any = changed;
    npairs_out = npairs;
}

// First force sweep of a step over every alive particle (also the only one for the
// jacobi order and for the kick-less t = 0 pass).
template <int SEL, bool ADIA, bool UM, bool MAPPED>
__global__ void __launch_bounds__(kTile, 4) k_force_first(SweepArgs A) {
    Ctl *ctl = A.ctl;
    const int lo = A.pad - ctl->GL, hi = A.pad + A.na + ctl->GR;
    const double Hb = __longlong_as_double((long long)ctl->hmax_bits);
    bool changed;
    int npairs;
    if (blockIdx.x == 0 && threadIdx.x == 0) { ctl->epoch_cur = A.epoch; ctl->sweeps_used = 1; }
    const int tile = MAPPED ? (int)blockIdx.x + A.tile_base : (int)blockIdx.x;
    force_tile<SEL, ADIA, UM, true>(A, tile, lo, hi, Hb, A.epoch, changed, npairs);
    // pair statistics for the roofline accounting (one atomic per warp)
    npairs = __reduce_add_sync(0xffffffffu, npairs);
    if ((threadIdx.x & 31) == 0 && npairs)
        atomicAdd(&ctl->force_pairs[(blockIdx.x * (kTile / 32) + (threadIdx.x >> 5)) % kStatBins], (unsigned long long)npairs);
}

// Refinement of the ordered sweep: ONE persistent cooperative launch that repeats
// "recompute every tile in range of a particle whose kicked (v,u) changed in the
// previous sweep" until a whole sweep changes nothing.  The dependency of particle i on
// the fully-kicked (v,u) of its lower-index neighbours is strictly lower triangular, so
// this fixed point exists, is unique, is reached after finitely many sweeps and IS the
// result of the reference's sequential in-place sweep.
// A sweep works through the work list its predecessor filled (tiles are appended once per
// sweep, guarded by the atomicMax on their stamp); a sweep that appends nothing proves
// that every particle was last computed from the final values.
template <bool ADIA, bool UM>
__global__ void __launch_bounds__(kTile, 3) k_force_refine(SweepArgs A, int max_sweeps) {
    cg::grid_group grid = cg::this_grid();
    Ctl *ctl = A.ctl;
    const int lo = A.pad - ctl->GL, hi = A.pad + A.na + ctl->GR;
    const double Hb = __longlong_as_double((long long)ctl->hmax_bits);
    int epoch = ctl->epoch_cur;   // stamp of the latest sweep, which filled wl[(epoch + 1) % 3]
    int used = ctl->sweeps_used;
    int peak = 0;
    grid.sync();                  // everybody has read the hand-over words before block 0 rewrites them
    // Three rotating lists make ONE grid-wide barrier per sweep enough: sweep e reads list e % 3, fills
    // (e + 1) % 3 and clears (e + 2) % 3, which every CTA finished reading before the previous barrier.
    while (true) {
        const int cur = (epoch + 1) % 3;
        const int nwork = *((volatile int *)&ctl->nwork[cur]);
        if (nwork == 0) break;
        if (used >= max_sweeps) {
            if (blockIdx.x == 0 && threadIdx.x == 0) raise(ctl, E_NOT_CONVERGED, -1);
            break;
        }
        epoch++;
        used++;
        peak = max(peak, nwork);
        if (blockIdx.x == 0 && threadIdx.x == 0) ctl->nwork[(epoch + 2) % 3] = 0;
        for (int w = blockIdx.x; w < nwork; w += gridDim.x) {
            const int tile = __ldcg(A.wl[cur] + w);
            bool changed;
            int npairs;
            force_tile<SEL_INDEX, ADIA, UM, false>(A, tile, lo, hi, Hb, epoch, changed, npairs);
        }
        grid.sync();
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && peak > ctl->nwork_peak) ctl->nwork_peak = peak;
    if (blockIdx.x == 0 && threadIdx.x == 0) { ctl->sweeps_used = used; ctl->epoch_cur = epoch; }
}

// ------------------------------------------------------------------------------------
// The ordered sweep in ONE pass (k_force_sweep).
//
// The reference's second loop (sph_simulation.cpp:94-107) is a recurrence along the array index: particle i needs
// the fully kicked (v, u) of its lower-index neighbours.  With the array kept x-sorted and the index order equal
// to the x order (always, except for the handful of particles that have overtaken a neighbour), "lower index"
// means "to the left", a few slots away.  The particle set is cut into segments; a GROUP of G lanes walks one
// segment from left to right in chunks of G particles:
//
//   phase P (lanes in parallel, lane = particle): the right-hand neighbours (seen in their old state, known),
//           the left-hand neighbours in earlier chunks (final values known), and for the left-hand neighbours
//           INSIDE the chunk everything that does not depend on their kicked values (pair_geom);
//   phase S (the chain): for t = 0 .. G-1 lane t's sums are complete, it is kicked, its (v, u) are broadcast,
//           and the lanes to its right that see it add the pair (pair_apply, a dozen operations).
//
// A segment does not know the final values of the particles to its left.  It starts B particles early
// ("burn-in") from nothing and discards what it computes there: the influence of a wrong left boundary decays
// by an order of magnitude per neighbour hop, so after a few dozen particles the values are the exact ones.
// "Are" is checked, not assumed: the burn-in results for the last G-1 particles before the segment are kept
// (K.scr) and compared BIT FOR BIT with what the left segment stored for them (k_sweep_verify).  By induction
// over the index every stored value is then the sequential sweep's.  Whatever this pass cannot settle —
// a failed comparison, a neighbourhood wider than the G-1 slots a group looks back / ahead (judged with the
// largest h of the surrounding tiles, `thm`), a right-hand neighbour with a lower index — queues the tile for
// the work-list refinement (k_force_refine), which re-evaluates it from the stored values, propagates any change,
// and reaches the same fixed point.  Both kernels evaluate a particle with the same functions in the same
// order, so an unflagged particle passes that evaluation unchanged.
constexpr int kSweepThreads = 128;
#ifndef SPHB_SWEEP_V
#define SPHB_SWEEP_V 1    // build-time experiments: 1 = prefetch during the chain (measured: -5 %), 2 = two right-hand pairs per round (slower: spills)
#endif

struct SerialArgs {
    SweepArgs S;
    const unsigned *thm;   // per tile of kTile slots: largest h (float bits, rounded up), from the density pass
    int nthm;              // tiles in thm
    double *scr;           // [nseg][2][G-1]: burn-in results next to the segment start, for k_sweep_verify
    int L, B, nseg;        // segment length, burn-in length (multiples of G), number of segments
};

template <int G>
struct SweepGrp {          // shared memory of one group: three chunks of particle records + the pair constants of a chunk
    static constexpr int W = 3 * G, NK = G * (G - 1) / 2;
    double2 hcA[W], rkA[W], hcB[W], rkB[W];   // (h, c5), (rho, kap) of the new (A) and the old (B) state
    double x[W], vB[W], uB[W], vA[W], uA[W];
    double kc[4][NK];      // PairK of lane l's pair with lane t < l at [t (G-1) - t (t-1)/2 + l - t - 1]
    unsigned idx[W];
};

__device__ __forceinline__ void queue_tile(const SweepArgs &A, int t, int epoch) {
    if (atomicMax(A.dirty + t, epoch) < epoch) {
        const int slot = atomicAdd(&A.ctl->nwork[(epoch + 1) % 3], 1);
        A.wl[(epoch + 1) % 3][slot] = t;
    }
}

template <bool ADIA, int G>
__global__ void __launch_bounds__(kSweepThreads, G == 8 ? 4 : 3) k_force_sweep(SerialArgs K) {
    constexpr int R = G - 1;
    using Grp = SweepGrp<G>;
    extern __shared__ __align__(16) unsigned char sweep_smem[];
    const SweepArgs &A = K.S;
    Ctl *ctl = A.ctl;
    const int gl = threadIdx.x % G;                  // lane within the group
    const int grp = threadIdx.x / G;
    const int seg = blockIdx.x * (kSweepThreads / G) + grp;
    if (blockIdx.x == 0 && threadIdx.x == 0) { ctl->epoch_cur = A.epoch; ctl->sweeps_used = 1; }
    if (seg >= K.nseg) return;                       // whole groups leave together
    Grp &S = reinterpret_cast<Grp *>(sweep_smem)[grp];
    const unsigned gmask = ((G == 32) ? 0xffffffffu : ((1u << G) - 1u)) << (((threadIdx.x & 31) / G) * G);
    const int lo = A.pad - ctl->GL, hi = A.pad + A.na + ctl->GR;
    const double Hb = __longlong_as_double((long long)ctl->hmax_bits);
    const double reach_glob = kRadius * Hb;
    const int seg_lo = A.pad + seg * K.L, seg_hi = min(seg_lo + K.L, A.pad + A.na);
    const bool halo_l0 = ctl->halo_left != 0;
    // first slot this group looks at.  Segment 0 starts at the wall ghosts; on a slab edge its burn-in is the
    // neighbouring slab's halo, of which twice the usual burn-in is walked (not all of it: the launch is one wave of
    // equally long segments there, and this group would finish last)
    const int st = (seg == 0) ? (halo_l0 ? max(lo, A.pad - 2 * K.B) : lo) : max(lo, seg_lo - K.B);
    const int nch = (seg_hi - st + G - 1) / G;
    const bool halo_l = ctl->halo_left != 0, halo_r = ctl->halo_right != 0;

    struct Pre { double x, vB, uB, gs, pst; double2 hcA, rkA, hcB, rkB; unsigned idx; };
    auto fetch = [&](int c) -> Pre {
        Pre p;
        const int s = st + c * G + gl;
        if (s >= lo && s < hi) {
            p.x = A.F.x[s]; p.idx = A.F.idx[s];
            const double2 *r = reinterpret_cast<const double2 *>(A.F.rec + s);
            p.hcA = r[0]; p.rkA = r[1]; p.hcB = r[2]; p.rkB = r[3];
            p.vB = A.F.vB[s]; p.uB = A.F.uB[s];
            p.gs = A.F.gsum[s]; p.pst = A.Pstale[s];
        } else {
            p.x = s < lo ? -INFINITY : INFINITY; p.idx = 0xffffffffu;
            p.hcA = p.rkA = p.hcB = p.rkB = make_double2(0.0, 0.0);
            p.vB = p.uB = p.gs = p.pst = 0.0;
        }
        return p;
    };
    auto put = [&](int buf, const Pre &p) {
        const int q = buf * G + gl;
        S.x[q] = p.x; S.idx[q] = p.idx; S.hcA[q] = p.hcA; S.rkA[q] = p.rkA; S.hcB[q] = p.hcB; S.rkB[q] = p.rkB;
        S.vB[q] = p.vB; S.uB[q] = p.uB; S.vA[q] = p.vB; S.uA[q] = p.uB;
    };
    double gs_c, pst_c, gs_n, pst_n;
    {
        const Pre p0 = fetch(0), p1 = fetch(1);
        put(0, p0); put(1, p1);
        Pre none;                                    // "chunk -1": nothing to the left of the start
        none.x = -INFINITY; none.idx = 0xffffffffu;
        none.hcA = none.rkA = none.hcB = none.rkB = make_double2(0.0, 0.0);
        none.vB = none.uB = none.gs = none.pst = 0.0;
        put(2, none);
        gs_c = p0.gs; pst_c = p0.pst; gs_n = p1.gs; pst_n = p1.pst;
    }
    __syncwarp(gmask);
    int npairs_tot = 0, nmoved = 0, nflag = 0;
    int tl_seen = -1;
    double reach_loc = 0.0, x_far_l = -INFINITY, x_far_r = INFINITY;

#pragma unroll 1
    for (int c = 0; c < nch; c++) {
        const int bc = c % 3, bp = (c + 2) % 3, bn = (c + 1) % 3;
#if !(SPHB_SWEEP_V & 1)
        const Pre nx = fetch(c + 2);                 // replaces chunk c - 1 once phase P is done with it
#endif
        const int s = st + c * G + gl;
        const int me = bc * G + gl;
        const unsigned idxi = S.idx[me];
        const bool eval = s < seg_hi && !(idxi & kGhostBit);   // alive (own or a neighbouring slab's); free slots carry 0xffffffff
        const bool store = eval && s >= seg_lo;
        const double vi = S.vB[me], ui = S.uB[me];
        Own<ADIA> o;
        {
            Rec ro;
            ro.h = S.hcA[me].x; ro.c5 = S.hcA[me].y; ro.rho = S.rkA[me].x; ro.kap = S.rkA[me].y;
            o.set(S.x[me], idxi, ro, vi, ui, pst_c);
        }
        double aL = 0.0, eL = 0.0, aR = 0.0, eR = 0.0;
        unsigned vmask = 0u, dmask = 0u;
        bool flag = false;
        if (eval) {
            // the scan bound: 2.5 x the largest h of the three tiles around this particle, never above the global one;
            // looked up again only when the particle enters another tile
            const int tl = s / kTile;
            if (tl != tl_seen) {
                tl_seen = tl;
                unsigned hb = K.thm[tl];
                if (tl > 0) hb = max(hb, K.thm[tl - 1]);
                if (tl + 1 < K.nthm) hb = max(hb, K.thm[tl + 1]);
                reach_loc = fmin(reach_glob, kRadius * (double)__uint_as_float(hb));
                // tiles further away than the neighbouring ones are only covered by the global bound: the first slot
                // of tile tl - 1 and the last of tile tl + 1 must be out of its reach
                const int fl = (tl - 1) * kTile, fr = (tl + 2) * kTile - 1;
                x_far_l = fl > lo ? __ldg(A.F.x + fl) : -INFINITY;
                x_far_r = fr < hi - 1 ? __ldg(A.F.x + fr) : INFINITY;
            }
            // slot s + q, -G <= q < 2G, in the ring (the three chunks follow each other cyclically)
            auto rpos = [&](int q) { int p = me + q; p -= p >= 3 * G ? 3 * G : 0; p += p < 0 ? 3 * G : 0; return p; };
            int dL, dR;                              // candidates per side; G = more than a group can hold
            if (G <= 8) {
                // few slots: every one of them is a candidate (pair_geom drops those out of reach at once); only
                // whether the first slot beyond the group's window could still interact has to be looked at
                dL = (o.x - S.x[rpos(-G)]) < reach_loc ? G : R;
                dR = (S.x[rpos(G)] - o.x) < reach_loc ? G : R;
            } else {
                dL = 0; dR = 0;
#pragma unroll 1
                while (dL < G && (o.x - S.x[rpos(-(dL + 1))]) < reach_loc) dL++;
#pragma unroll 1
                while (dR < G && (S.x[rpos(dR + 1)] - o.x) < reach_loc) dR++;
            }
            if (store) {
                if (dL == G || dR == G) flag = true;
                // the array ends inside the window and its last particle is within reach: next to a wall that is
                // the end of the world, on a slab edge the halo is thinner than the kernel reaches
                if ((halo_l && s - G < lo && (o.x - S.x[rpos(lo - s)]) < reach_loc) ||
                    (halo_r && s + G >= hi && (S.x[rpos(hi - 1 - s)] - o.x) < reach_loc)) raise(ctl, E_HALO, s);
                if ((o.x - x_far_l) < reach_glob || (x_far_r - o.x) < reach_glob) flag = true;
            }
            dL = min(dL, R); dR = min(dR, R);
            npairs_tot++;                            // self: its energy term is evaluated by the reference too (and is exactly 0)
            // left-hand side, farthest first
#pragma unroll 1
            for (int d = dL; d >= 1; d--) {
                const int p = gl - d;
                const int rp = rpos(-d);
                const bool useA = S.idx[rp] < o.idx;             // the reference's sweep has already updated j
                const double2 hc = useA ? S.hcA[rp] : S.hcB[rp], rk = useA ? S.rkA[rp] : S.rkB[rp];
                Rec rj;
                rj.h = hc.x; rj.c5 = hc.y; rj.rho = rk.x; rj.kap = rk.y;
                PairK pk;
                if (!pair_geom<ADIA>(o, o.x - S.x[rp], 1.0, rj, pk)) continue;
                npairs_tot++;
                if (p < 0) {                                     // an earlier chunk: its final values are in the ring
                    pair_apply<ADIA, true>(o, pk, useA ? S.vA[rp] : S.vB[rp], useA ? S.uA[rp] : S.uB[rp], 0.0, aL, eL);
                } else {
                    const int q = p * R - (p * (p - 1)) / 2 + (gl - p - 1);
                    if (useA) {                                  // this chunk: added when lane p has been kicked
                        S.kc[0][q] = pk.K2; S.kc[1][q] = pk.irho; S.kc[2][q] = pk.gb; S.kc[3][q] = pk.G1;
                        vmask |= 1u << p;
                    } else {                                     // a ghost, or a higher index: seen in its old state
                        double ta = 0.0, te = 0.0;
                        pair_apply<ADIA, true>(o, pk, S.vB[rp], S.uB[rp], 0.0, ta, te);
                        S.kc[0][q] = ta; S.kc[1][q] = te;
                        dmask |= 1u << p;
                    }
                }
            }
            // right-hand side, farthest first: old state
#if (SPHB_SWEEP_V & 2)
            // two neighbours per round: their evaluations are independent and the compiler can overlap the two dependent
            // FP64 chains (the sums still receive the farther one first)
            int d = dR;
#pragma unroll 1
            for (; d >= 2; d -= 2) {
                const int rp0 = rpos(d), rp1 = rpos(d - 1);
                if (S.idx[rp0] < o.idx || S.idx[rp1] < o.idx) flag = true;
                const double2 hc0 = S.hcB[rp0], rk0 = S.rkB[rp0], hc1 = S.hcB[rp1], rk1 = S.rkB[rp1];
                Rec r0, r1;
                r0.h = hc0.x; r0.c5 = hc0.y; r0.rho = rk0.x; r0.kap = rk0.y;
                r1.h = hc1.x; r1.c5 = hc1.y; r1.rho = rk1.x; r1.kap = rk1.y;
                PairK k0, k1;
                const bool ok0 = pair_geom<ADIA>(o, S.x[rp0] - o.x, -1.0, r0, k0);
                const bool ok1 = pair_geom<ADIA>(o, S.x[rp1] - o.x, -1.0, r1, k1);
                if (ok0) { pair_apply<ADIA, true>(o, k0, S.vB[rp0], S.uB[rp0], 0.0, aR, eR); npairs_tot++; }
                if (ok1) { pair_apply<ADIA, true>(o, k1, S.vB[rp1], S.uB[rp1], 0.0, aR, eR); npairs_tot++; }
            }
            if (d == 1) {
                const int rp = rpos(1);
                if (S.idx[rp] < o.idx) flag = true;
                const double2 hc = S.hcB[rp], rk = S.rkB[rp];
                Rec rj;
                rj.h = hc.x; rj.c5 = hc.y; rj.rho = rk.x; rj.kap = rk.y;
                PairK pk;
                if (pair_geom<ADIA>(o, S.x[rp] - o.x, -1.0, rj, pk)) {
                    pair_apply<ADIA, true>(o, pk, S.vB[rp], S.uB[rp], 0.0, aR, eR);
                    npairs_tot++;
                }
            }
#else
#pragma unroll 1
            for (int d = dR; d >= 1; d--) {
                const int rp = rpos(d);
                if (S.idx[rp] < o.idx) flag = true;              // overtaken: it is seen in its NEW state, which is not known yet
                const double2 hc = S.hcB[rp], rk = S.rkB[rp];
                Rec rj;
                rj.h = hc.x; rj.c5 = hc.y; rj.rho = rk.x; rj.kap = rk.y;
                PairK pk;
                if (pair_geom<ADIA>(o, S.x[rp] - o.x, -1.0, rj, pk)) {
                    pair_apply<ADIA, true>(o, pk, S.vB[rp], S.uB[rp], 0.0, aR, eR);
                    npairs_tot++;
                }
            }
#endif
        }
        __syncwarp(gmask);                           // everybody is done with chunk c - 1
#if (SPHB_SWEEP_V & 1)
        // the loads of chunk c + 2 fly while the chain runs (its registers are not live during phase P)
        const Pre nx = fetch(c + 2);
#else
        {
            const int q = bp * G + gl;
            S.x[q] = nx.x; S.idx[q] = nx.idx; S.hcA[q] = nx.hcA; S.rkA[q] = nx.rkA; S.hcB[q] = nx.hcB; S.rkB[q] = nx.rkB;
            S.vB[q] = nx.vB; S.uB[q] = nx.uB; S.vA[q] = nx.vB; S.uA[q] = nx.uB;
        }
#endif
        // the chain
#pragma unroll
        for (int t = 0; t < G - 1; t++) {
            // merged exactly as the other force kernels merge them: (0 - left) + right, (0 + left) - right
            double a_s = (0.0 - aL) + aR, e_s = (0.0 + eL) - eR;
            a_s *= A.m0; e_s *= A.m0;
            a_s = fma(-o.PrRho, gs_c, a_s);
            // second half-kick, sph_simulation.cpp:105-106, operation by operation
            const double vf = __dadd_rn(vi, __dmul_rn(a_s, A.half_dt));
            const double uf = __dadd_rn(ui, __dmul_rn(e_s, A.half_dt));
            const double bv = __shfl_sync(gmask, vf, t, G), bu = __shfl_sync(gmask, uf, t, G);
            if (gl > t) {
                const int q = t * R - (t * (t - 1)) / 2 + (gl - t - 1);
                if ((vmask >> t) & 1u) {
                    PairK pk;
                    pk.K2 = S.kc[0][q]; pk.irho = S.kc[1][q]; pk.gb = S.kc[2][q]; pk.G1 = S.kc[3][q];
                    pair_apply<ADIA, true>(o, pk, bv, bu, 0.0, aL, eL);
                } else if ((dmask >> t) & 1u) {
                    aL += S.kc[0][q]; eL += S.kc[1][q];
                }
            }
        }
#if (SPHB_SWEEP_V & 1)
        {
            const int q = bp * G + gl;
            S.x[q] = nx.x; S.idx[q] = nx.idx; S.hcA[q] = nx.hcA; S.rkA[q] = nx.rkA; S.hcB[q] = nx.hcB; S.rkB[q] = nx.rkB;
            S.vB[q] = nx.vB; S.uB[q] = nx.uB; S.vA[q] = nx.vB; S.uA[q] = nx.uB;
        }
#endif
        // every lane's left-hand sums are complete now (lane t's were when it was broadcast: same inputs, same bits)
        double my_a = (0.0 - aL) + aR, my_e = (0.0 + eL) - eR;
        my_a *= A.m0; my_e *= A.m0;
        my_a = fma(-o.PrRho, gs_c, my_a);
        const double my_vf = __dadd_rn(vi, __dmul_rn(my_a, A.half_dt));
        const double my_uf = __dadd_rn(ui, __dmul_rn(my_e, A.half_dt));
        if (eval) { S.vA[me] = my_vf; S.uA[me] = my_uf; }
        bool moved = false;
        if (store) {
            A.a[s] = my_a; A.du[s] = my_e; A.Pout[s] = o.P;
            A.vn[s] = my_vf; A.un[s] = my_uf;
            moved = __double_as_longlong(my_vf) != __double_as_longlong(vi) || __double_as_longlong(my_uf) != __double_as_longlong(ui);
            if (flag) { queue_tile(A, (s - A.pad) / kTile, A.epoch); nflag++; }
        } else if (eval) {
            if (s < A.pad) {
                // a neighbouring slab's particle: my estimate of its final values; the owner's arrive with the
                // ITERATE message and are compared with these
                A.vn[s] = my_vf; A.un[s] = my_uf;
            } else if (s >= seg_lo - R) {
                K.scr[(size_t)(2 * seg) * R + (s - (seg_lo - R))] = my_vf;
                K.scr[(size_t)(2 * seg + 1) * R + (s - (seg_lo - R))] = my_uf;
            }
        }
        if ((__ballot_sync(gmask, moved) & gmask) && gl == 0) nmoved++;
        gs_c = gs_n; pst_c = pst_n; gs_n = nx.gs; pst_n = nx.pst;
        __syncwarp(gmask);
    }
    // statistics: pairs for the roofline accounting, chunks with a non-zero kick, flagged particles
    npairs_tot = __reduce_add_sync(gmask, npairs_tot);
    nflag = __reduce_add_sync(gmask, nflag);
    if (gl == 0) {
        atomicAdd(&ctl->force_pairs[seg % kStatBins], (unsigned long long)npairs_tot);
        if (nmoved) atomicAdd(&ctl->sweep_moved_chunks, nmoved);
        if (nflag) atomicAdd(&ctl->sweep_flagged, nflag);
    }
}

// Burn-in check of the serial sweep: what segment `seg` computed for the R particles before its first one must
// be, bit for bit, what the segment to its left stored for them.  If not, the segment's first tile is
// re-evaluated by the work-list refinement (which passes any change on to the right).
__global__ void k_sweep_verify(SerialArgs K, int R) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int seg = 1 + t / R, r = t % R;
    if (seg >= K.nseg) return;
    const SweepArgs &A = K.S;
    const int seg_lo = A.pad + seg * K.L;
    const int s = seg_lo - R + r;
    const double v = K.scr[(size_t)(2 * seg) * R + r], u = K.scr[(size_t)(2 * seg + 1) * R + r];
    if (__double_as_longlong(v) != __double_as_longlong(A.vn[s]) || __double_as_longlong(u) != __double_as_longlong(A.un[s])) {
        atomicAdd(&A.ctl->sweep_verify_fail, 1);
        queue_tile(A, (seg_lo - A.pad) / kTile, A.ctl->epoch_cur);
    }
}

// ------------------------------------------------------------------------------------
// First half of the step (sph_simulation.cpp:59-73): v += a dt/2; u += du dt/2;
// x += v dt, each product and sum rounded as written.  Fused with it: the ghost
// predicate of ghost_particles.cpp:45-55 (stream compaction by atomic append: only the
// handful of particles next to a wall ever qualify) and the x-order check.
struct KickArgs {
    const double *x;   // positions before the drift
    double *xo;        // positions after the drift (a different buffer: the order check re-derives
                       // the left neighbour's new position from the OLD arrays, race-free)
    const double *v, *u, *a, *du, *h;
    double *vh, *uh;
    int *cand;
    Ctl *ctl;
    int pad, na, cand_cap;
    double dt, half_dt, limit;
    int do_kick;      // 0: ghost selection + order check only (positions unchanged, xo unused)
};

__device__ __forceinline__ bool ghost_predicate(double pos, double h, double limit) {
    const double reach = __dmul_rn(kRadius, h);
    const double max_pos = __dsub_rn(limit, reach);
    const double min_pos = __dadd_rn(-limit, reach);
    return pos > max_pos || pos < min_pos;
}

__global__ void __launch_bounds__(kTile) k_kick_drift(KickArgs A) {
    __shared__ double sx[kTile + 1];
    const int g = blockIdx.x * kTile + threadIdx.x;
    const int s = A.pad + g;
    double xn = 0.0;
    if (g < A.na) {
        if (A.do_kick) {
            const double vhalf = __dadd_rn(A.v[s], __dmul_rn(A.a[s], A.half_dt));
            const double uhalf = __dadd_rn(A.u[s], __dmul_rn(A.du[s], A.half_dt));
            xn = __dadd_rn(A.x[s], __dmul_rn(vhalf, A.dt));
            A.vh[s] = vhalf;
            A.uh[s] = uhalf;
            A.xo[s] = xn;
        } else {
            xn = A.x[s];
        }
        const double hh = A.h[s];
        if (hh < kEps) raise(A.ctl, E_GHOST_ZERO_H, s);
        if (ghost_predicate(xn, hh, A.limit)) {
            int k = atomicAdd(&A.ctl->ncand, 1);
            if (k < A.cand_cap) A.cand[k] = s;
        }
        sx[threadIdx.x + 1] = xn;
    }
    if (threadIdx.x == 0) {
        double xp = -INFINITY;
        if (g > 0 && g < A.na) {
            if (A.do_kick) {
                const double vp = __dadd_rn(A.v[s - 1], __dmul_rn(A.a[s - 1], A.half_dt));
                xp = __dadd_rn(A.x[s - 1], __dmul_rn(vp, A.dt));
            } else xp = A.x[s - 1];
        }
        sx[0] = xp;
    }
    __syncthreads();
    if (g < A.na && sx[threadIdx.x] > xn) A.ctl->unsorted = 1;
}

__global__ void __launch_bounds__(kTile) k_check_sorted(const double *x, Ctl *ctl, int pad, int na) {
    const int g = blockIdx.x * kTile + threadIdx.x;
    if (g > 0 && g < na && x[pad + g - 1] > x[pad + g]) ctl->unsorted = 1;
}

// ------------------------------------------------------------------------------------
// Ghost set-up (ghost_particles.cpp:57-115) for the selected candidates: mirrored
// position, negated velocity, everything else copied; placed x-sorted into the pads and
// remembered in ascending source-index order for the download (the reference appends
// them in alive-index order).  One CTA: the candidate count is O(10) at any N.
struct GhostArgs {
    double *x, *vd, *ud, *a, *du, *h, *rho, *P, *m;   // vd/ud: the v,u arrays the ghosts live in
    double *ad, *dud;                                  // the acc, du_dt arrays the ghosts live in (a, du: sources)
    const double *vs, *us;                             // the v,u arrays the sources are read from
    unsigned *idx;
    const int *cand;
    // staging (capacity 2*pad each)
    double *gx, *gv, *gu, *ga, *gdu, *gh, *grho, *gP, *gm;
    unsigned *gkey;
    int *gorder;       // download order k -> pad slot
    Ctl *ctl;
    int pad, na, cand_cap;
    double limit;
    int from_alive;    // 1: stage from the alive candidates; 0: staging already filled (upload)
    int nstaged;       // number staged when from_alive = 0
};

__global__ void __launch_bounds__(1024) k_ghost_build(GhostArgs G) {
    Ctl *ctl = G.ctl;
    __shared__ int s_nl;
    int nc = G.from_alive ? ctl->ncand : G.nstaged;
    if (nc > G.cand_cap) {
        if (threadIdx.x == 0) { raise(ctl, E_CAPACITY, -1); ctl->GL = 0; ctl->GR = 0; }
        return;
    }
    if (G.from_alive) {
        for (int c = threadIdx.x; c < nc; c += blockDim.x) {
            const int s = G.cand[c];
            const double p = G.x[s];
            // ghost_particles.cpp:60-69: the side is chosen by the sign of x
            const double vec = (p < 0) ? __dsub_rn(-G.limit, p) : __dsub_rn(G.limit, p);
            G.gx[c] = __dadd_rn(p, __dmul_rn(2.0, vec));
            G.gv[c] = -G.vs[s];
            G.gu[c] = G.us[s];
            G.ga[c] = G.a[s];
            G.gdu[c] = G.du[s];
            G.gh[c] = G.h[s];
            G.grho[c] = G.rho[s];
            G.gP[c] = G.P[s];
            G.gm[c] = G.m[s];
            G.gkey[c] = G.idx[s];
        }
    }
    if (threadIdx.x == 0) s_nl = 0;
    __syncthreads();
    int nl_local = 0;
    for (int c = threadIdx.x; c < nc; c += blockDim.x) nl_local += (G.gx[c] < 0) ? 1 : 0;
    if (nl_local) atomicAdd(&s_nl, nl_local);
    __syncthreads();
    const int GL = s_nl, GR = nc - GL;
    if (GL > G.pad || GR > G.pad) {
        if (threadIdx.x == 0) { raise(ctl, E_CAPACITY, -1); ctl->GL = 0; ctl->GR = 0; }
        return;
    }
    for (int c = threadIdx.x; c < nc; c += blockDim.x) {
        const double xc = G.gx[c];
        const bool left = xc < 0;
        const unsigned kc = G.gkey[c];
        int rank_x = 0, rank_k = 0;
        for (int d = 0; d < nc; d++) {
            const double xd = G.gx[d];
            if ((xd < 0) == left && (xd < xc || (xd == xc && d < c))) rank_x++;
            const unsigned kd = G.gkey[d];
            if (kd < kc || (kd == kc && d < c)) rank_k++;
        }
        const int dest = left ? (G.pad - GL + rank_x) : (G.pad + G.na + rank_x);
        G.x[dest] = xc;
        G.vd[dest] = G.gv[c];
        G.ud[dest] = G.gu[c];
        G.ad[dest] = G.ga[c];
        G.dud[dest] = G.gdu[c];
        G.h[dest] = G.gh[c];
        G.rho[dest] = G.grho[c];
        G.P[dest] = G.gP[c];
        G.m[dest] = G.gm[c];
        G.idx[dest] = kGhostBit | kc;
        G.gorder[rank_k] = dest;
    }
    if (threadIdx.x == 0) { ctl->GL = GL; ctl->GR = GR; }
    // Every scan assumes that [pad - GL, pad + na + GR) is x-sorted as a whole.  A particle outside the wall, or wide
    // enough to be mirrored about the far wall, would put its ghost INSIDE the alive range: the scans would then stop
    // early and drop pairs silently (the reference's all-pairs loops do not).  Refuse instead.
    __syncthreads();
    if (threadIdx.x == 0 && G.na > 0 &&
        ((GL > 0 && G.x[G.pad - 1] > G.x[G.pad]) || (GR > 0 && G.x[G.pad + G.na] < G.x[G.pad + G.na - 1])))
        raise(ctl, E_UNSORTED, GL > 0 && G.x[G.pad - 1] > G.x[G.pad] ? G.pad - 1 : G.pad + G.na);
}

// copy ghost (pad) entries of v,u between the half-kicked and the current arrays
__global__ void k_copy_pads(const double *vs, const double *us, double *vd, double *ud, const Ctl *ctl,
                            int pad, int na) {
    const int GL = ctl->GL, GR = ctl->GR;
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < GL + GR; k += gridDim.x * blockDim.x) {
        const int s = k < GL ? (pad - GL + k) : (pad + na + (k - GL));
        vd[s] = vs[s];
        ud[s] = us[s];
    }
}

// ------------------------------------------------------------------------------------
// AoS (the reference's Particle, 88 bytes) <-> SoA.
struct AosRec {
    int id, pad0;
    double mass, pos, vel, acc, h, du_dt, u, density, pressure;
    int type, pad1;
};
static_assert(sizeof(AosRec) == 88, "Particle layout");

struct PackArgs {
    double *x, *v, *u, *a, *du, *h, *rho, *P, *m;
    unsigned *idx;
    int *slot_id;
    double *gx, *gv, *gu, *ga, *gdu, *gh, *grho, *gP, *gm;
    unsigned *gkey;
    const double *vg, *ug;     // arrays holding the ghosts' v,u at download time
    const int *gorder;
    const int *cnt;
    AosRec *aos;
    int *out_newton, *out_nbr;
    const Ctl *ctl;
    int pad, na, ng;
    unsigned idx_base;   // global index of the first local particle (x-slabs), else 0
    int by_slot;         // x-slab runs: records travel in x order and carry their global index in `id`
    double m_first;      // upload: mass of record 0 (uniform masses are taken from a register by the kernels)
    Ctl *ctl_rw;
};

__global__ void k_unpack(PackArgs A) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= A.na + A.ng) return;
    const AosRec r = A.aos[i];
    A.slot_id[i] = r.id;
    if (r.mass != A.m_first) A.ctl_rw->mass_differs = 1;
    if (i < A.na) {
        const int s = A.pad + i;
        A.x[s] = r.pos; A.v[s] = r.vel; A.u[s] = r.u; A.a[s] = r.acc; A.du[s] = r.du_dt;
        A.h[s] = r.h; A.rho[s] = r.density; A.P[s] = r.pressure; A.m[s] = r.mass;
        A.idx[s] = A.by_slot ? (unsigned)r.id : (unsigned)i + A.idx_base;
    } else {
        const int k = i - A.na;
        A.gx[k] = r.pos; A.gv[k] = r.vel; A.gu[k] = r.u; A.ga[k] = r.acc; A.gdu[k] = r.du_dt;
        A.gh[k] = r.h; A.grho[k] = r.density; A.gP[k] = r.pressure; A.gm[k] = r.mass;
        A.gkey[k] = (unsigned)k;
    }
}

__global__ void k_pack(PackArgs A) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int ng = A.ng;
    if (t >= A.na + ng) return;
    int s, out;
    const double *vv, *uu;
    int type;
    if (t < A.na) {
        s = A.pad + t;
        out = A.by_slot ? t : (int)(A.idx[s] - A.idx_base);
        vv = A.v; uu = A.u; type = 0;
    } else {
        s = A.gorder[t - A.na];
        out = t;
        vv = A.vg; uu = A.ug; type = 1;
    }
    AosRec r;
    r.id = (A.by_slot && t < A.na) ? (int)A.idx[s] : A.slot_id[out]; r.pad0 = 0;
    r.mass = A.m[s]; r.pos = A.x[s]; r.vel = vv[s]; r.acc = A.a[s]; r.h = A.h[s];
    r.du_dt = A.du[s]; r.u = uu[s]; r.density = A.rho[s]; r.pressure = A.P[s];
    r.type = type; r.pad1 = 0;
    A.aos[out] = r;
    if (A.out_newton) {
        const int c = A.cnt[s];
        A.out_newton[out] = (c >> 24) & 0xff;
        A.out_nbr[out] = c & 0xffffff;
    }
}

// one double per particle, sorted slot -> array-index order (same mapping as k_pack)
__global__ void k_pack_f64(const double *src, double *out, const unsigned *idx, const int *gorder, const Ctl *ctl,
                           int pad, int na, int ng, unsigned idx_base, int by_slot) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= na + ng) return;
    if (t < na) {
        const int s = pad + t;
        out[by_slot ? t : (int)(idx[s] - idx_base)] = src[s];
    } else {
        out[t] = src[gorder[t - na]];
    }
}

// ------------------------------------------------------------------------------------
// x-slab decomposition: the K particles next to a slab edge travel to the
// neighbouring rank and fill ITS pad on the facing side, in the same ascending-x order, so that
// every kernel above runs unchanged.  Messages are rows of K doubles plus a header row.
//   BASE    after the drift:  x, m, v_half, u_half, h_old, rho_old, P, idx            (8 rows)
//   RECORDS after the density pass: the owner's Rec2 and own-h gradient sum            (9 rows)
//   ITERATE during the ordered sweep: current fully kicked v, u                        (2 rows)
// header[0] = number of particles in the message.
enum { MSG_BASE = 0, MSG_RECORDS = 1, MSG_ITERATE = 2 };
__host__ __device__ inline int msg_rows(int what) { return what == MSG_ITERATE ? 2 : (what == MSG_RECORDS ? 9 : 8); }

struct SlabArgs {
    double *x, *m, *vh, *uh, *vn, *un, *h, *rho, *P;
    unsigned *idx;
    Rec2 *rec;
    double *gsum;    // own-h gradient sums (the serial sweep evaluates the left halo's particles as its burn-in)
    unsigned *thm;   // per-tile h maximum (see DensityArgs)
    int *dirty;
    int *wl[3];
    Ctl *ctl;
    int pad, na, K;
    int side;        // -1: my left edge / my left pad, +1: right
};

// my edge particles -> message (side = which of my edges)
__device__ __forceinline__ void slab_pack_body(const SlabArgs &A, int what, double *buf) {
    const int cnt = min(A.K, A.na);
    const int first = A.side < 0 ? A.pad : A.pad + A.na - cnt;
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < cnt; k += gridDim.x * blockDim.x) {
        const int s = first + k;
        if (what == MSG_BASE) {
            buf[0 * A.K + k] = A.x[s]; buf[1 * A.K + k] = A.m[s]; buf[2 * A.K + k] = A.vh[s];
            buf[3 * A.K + k] = A.uh[s]; buf[4 * A.K + k] = A.h[s]; buf[5 * A.K + k] = A.rho[s];
            buf[6 * A.K + k] = A.P[s]; buf[7 * A.K + k] = (double)A.idx[s];
        } else if (what == MSG_RECORDS) {
            const Rec2 r = A.rec[s];
            buf[0 * A.K + k] = r.A.h; buf[1 * A.K + k] = r.A.c5; buf[2 * A.K + k] = r.A.rho; buf[3 * A.K + k] = r.A.kap;
            buf[4 * A.K + k] = r.B.h; buf[5 * A.K + k] = r.B.c5; buf[6 * A.K + k] = r.B.rho; buf[7 * A.K + k] = r.B.kap;
            buf[8 * A.K + k] = A.gsum[s];
        } else {
            buf[0 * A.K + k] = A.vn[s]; buf[1 * A.K + k] = A.un[s];
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        buf[msg_rows(what) * A.K] = (double)cnt;
        if (what == MSG_ITERATE) A.ctl->halo_changed = 0;   // the unpack that follows this exchange sets it again
    }
}

__global__ void k_slab_pack(SlabArgs A, int what, double *buf) { slab_pack_body(A, what, buf); }
// both edges in one launch: blockIdx.y = 0 left, 1 right; a null buffer = no neighbour on that side
__global__ void k_slab_pack2(SlabArgs A, int what, double *buf_l, double *buf_r) {
    double *buf = blockIdx.y ? buf_r : buf_l;
    if (!buf) return;
    A.side = blockIdx.y ? 1 : -1;
    slab_pack_body(A, what, buf);
}

// message from the neighbour on `side` -> my pad on that side.  One CTA.
__device__ __forceinline__ void slab_unpack_body(const SlabArgs &A, int what, const double *buf) {
    Ctl *ctl = A.ctl;
    const int cnt = (int)buf[msg_rows(what) * A.K];
    if (cnt < 0 || cnt > A.K || cnt > A.pad) {
        if (threadIdx.x == 0) raise(ctl, E_CAPACITY, -1);
        return;
    }
    // a message from my LEFT neighbour holds ITS last particles: they end right before my first one
    const int first = A.side < 0 ? A.pad - cnt : A.pad + A.na;
    __shared__ int s_changed, s_tile;
    if (threadIdx.x == 0) { s_changed = 0; s_tile = A.side < 0 ? -1 : 0x7fffffff; }
    __syncthreads();
    const double Hb = __longlong_as_double((long long)ctl->hmax_bits);
    const int ntiles = (A.na + kTile - 1) / kTile;
    for (int k = threadIdx.x; k < cnt; k += blockDim.x) {
        const int s = first + k;
        if (what == MSG_BASE) {
            A.x[s] = buf[0 * A.K + k]; A.m[s] = buf[1 * A.K + k];
            const double v = buf[2 * A.K + k], u = buf[3 * A.K + k];
            A.vh[s] = v; A.uh[s] = u;
            A.vn[s] = v; A.un[s] = u;            // first guess of the neighbour's fully kicked values
            A.h[s] = buf[4 * A.K + k]; A.rho[s] = buf[5 * A.K + k]; A.P[s] = buf[6 * A.K + k];
            A.idx[s] = (unsigned)buf[7 * A.K + k];
        } else if (what == MSG_RECORDS) {
            Rec2 r;
            r.A.h = buf[0 * A.K + k]; r.A.c5 = buf[1 * A.K + k]; r.A.rho = buf[2 * A.K + k]; r.A.kap = buf[3 * A.K + k];
            r.B.h = buf[4 * A.K + k]; r.B.c5 = buf[5 * A.K + k]; r.B.rho = buf[6 * A.K + k]; r.B.kap = buf[7 * A.K + k];
            A.rec[s] = r;
            A.gsum[s] = buf[8 * A.K + k];
            // the scan bounds of the force pass must cover the neighbour's smoothing lengths too
            atomicMax(&ctl->hmax_bits, (unsigned long long)__double_as_longlong(fmax(r.A.h, r.B.h)));
            atomicMax(A.thm + s / kTile, __float_as_uint(__double2float_ru(fmax(r.A.h, r.B.h))));
        } else {
            const double v = buf[0 * A.K + k], u = buf[1 * A.K + k];
            if (__double_as_longlong(v) != __double_as_longlong(A.vn[s]) ||
                __double_as_longlong(u) != __double_as_longlong(A.un[s])) {
                A.vn[s] = v; A.un[s] = u;
                // Only particles that see this neighbour in its NEW state depend on the value: own
                // particles within reach whose array index is higher (the sweep reaches them later).
                // Without one, the new value is merely kept.
                const double xj = A.x[s];
                const unsigned idxj = A.idx[s];
                bool relevant = false;
                if (A.side < 0) {
                    int q = A.pad;
                    while (q < A.pad + A.na && A.x[q] - xj < kRadius * Hb) { relevant |= A.idx[q] > idxj; q++; }
                    if (relevant) atomicMax(&s_tile, (q - 1 - A.pad) / kTile);
                } else {
                    int q = A.pad + A.na - 1;
                    while (q >= A.pad && xj - A.x[q] < kRadius * Hb) { relevant |= A.idx[q] > idxj; q--; }
                    if (relevant) atomicMin(&s_tile, (q + 1 - A.pad) / kTile);
                }
                if (relevant) s_changed = 1;
            }
        }
    }
    __syncthreads();
    if (what == MSG_BASE && threadIdx.x == 0) {
        if (A.side < 0) { ctl->GL = cnt; ctl->halo_left = 1; } else { ctl->GR = cnt; ctl->halo_right = 1; }
    }
    if (what == MSG_BASE) {
        // The slabs must stay ordered ACROSS the edge too.  When the drift made the last particle of the left
        // slab overtake the first of the right one, both ranks hold the same 2K particles (their own edge and
        // the message) and do the same merge-split on them: the lower ones stay / go left, the upper ones
        // right, counts unchanged, no further message needed.  Everything a particle owns at this point of
        // the step is in the BASE message.
        __syncthreads();
        const int ko = min(A.K, A.na);
        const int lfirst = A.side < 0 ? A.pad - cnt : A.pad + A.na - ko;   // the left set of the 2K window
        const int nleft = A.side < 0 ? cnt : ko, nright = A.side < 0 ? ko : cnt;
        const int ntot = nleft + nright;
        __shared__ int s_inv;
        if (threadIdx.x == 0) {
            const int a = lfirst + nleft - 1, b = lfirst + nleft;         // last of the left set, first of the right set
            s_inv = (nleft > 0 && nright > 0 &&
                     (A.x[a] > A.x[b] || (A.x[a] == A.x[b] && A.idx[a] > A.idx[b]))) ? 1 : 0;
        }
        __syncthreads();
        if (s_inv) {
            if (ntot > 2 * kMergeMax) { if (threadIdx.x == 0) raise(ctl, E_UNSORTED, lfirst); return; }
            __shared__ double s_kx[2 * kMergeMax];
            __shared__ unsigned s_ki[2 * kMergeMax];
            for (int e = threadIdx.x; e < ntot; e += blockDim.x) { s_kx[e] = A.x[lfirst + e]; s_ki[e] = A.idx[lfirst + e]; }
            __syncthreads();
            // every thread carries up to two elements: read all their fields, rank them, write them back
            double f[2][9];
            unsigned fi[2];
            int dest[2] = {-1, -1};
            for (int q = 0; q < 2; q++) {
                const int e = threadIdx.x + q * blockDim.x;
                if (e >= ntot) continue;
                const int sl = lfirst + e;
                f[q][0] = A.x[sl]; f[q][1] = A.m[sl]; f[q][2] = A.vh[sl]; f[q][3] = A.uh[sl]; f[q][4] = A.vn[sl];
                f[q][5] = A.un[sl]; f[q][6] = A.h[sl]; f[q][7] = A.rho[sl]; f[q][8] = A.P[sl];
                fi[q] = A.idx[sl];
                int rank = 0;
                const double xe = s_kx[e];
                const unsigned ie = s_ki[e];
                for (int g = 0; g < ntot; g++) rank += (s_kx[g] < xe || (s_kx[g] == xe && s_ki[g] < ie)) ? 1 : 0;
                dest[q] = lfirst + rank;
                if ((e < nleft) != (rank < nleft)) atomicAdd(&ctl->migrated, 1);
            }
            __syncthreads();
            for (int q = 0; q < 2; q++) {
                if (dest[q] < 0) continue;
                const int sl = dest[q];
                A.x[sl] = f[q][0]; A.m[sl] = f[q][1]; A.vh[sl] = f[q][2]; A.uh[sl] = f[q][3]; A.vn[sl] = f[q][4];
                A.un[sl] = f[q][5]; A.h[sl] = f[q][6]; A.rho[sl] = f[q][7]; A.P[sl] = f[q][8];
                A.idx[sl] = fi[q];
            }
            __syncthreads();
            // deeper than the window?  then the own block is no longer sorted where the window begins / ends
            if (threadIdx.x == 0) {
                const int own0 = A.side < 0 ? A.pad : A.pad + A.na - ko;
                const bool bad = A.side < 0 ? (ko < A.na && A.x[own0 + ko - 1] > A.x[own0 + ko])
                                            : (ko < A.na && A.x[own0 - 1] > A.x[own0]);
                if (bad) raise(ctl, E_UNSORTED, own0);
            }
        }
        // how far the halo reaches from my edge particle: a smoothing length anywhere in the run whose support is
        // wider than this could interact across the halo unseen (checked with the all-reduced h maximum)
        __syncthreads();
        if (threadIdx.x == 0 && cnt > 0 && A.na > 0)
            ctl->halo_thick[A.side < 0 ? 0 : 1] = A.side < 0 ? (A.x[A.pad] - A.x[A.pad - cnt])
                                                              : (A.x[A.pad + A.na + cnt - 1] - A.x[A.pad + A.na - 1]);
    }
    if (what == MSG_ITERATE && s_changed) {
        // queue those tiles for the next sweep, exactly like a changed particle of my own would
        const int epoch = ctl->epoch_cur;
        const int ta = A.side < 0 ? 0 : min(s_tile, ntiles);
        const int tb = A.side < 0 ? min(s_tile, ntiles - 1) : ntiles - 1;
        for (int t = ta + threadIdx.x; t <= tb; t += blockDim.x) {
            if (atomicMax(A.dirty + t, epoch) < epoch) {
                const int slot = atomicAdd(&ctl->nwork[(epoch + 1) % 3], 1);
                A.wl[(epoch + 1) % 3][slot] = t;
            }
        }
        if (threadIdx.x == 0) ctl->halo_changed = 1;
    }
}

__global__ void __launch_bounds__(1024) k_slab_unpack(SlabArgs A, int what, const double *buf) {
    slab_unpack_body(A, what, buf);
}
// both pads in one launch: one CTA per side (blockIdx.x = 0 left, 1 right); a null buffer = no neighbour
__global__ void __launch_bounds__(1024) k_slab_unpack2(SlabArgs A, int what, const double *buf_l, const double *buf_r) {
    const double *buf = blockIdx.x ? buf_r : buf_l;
    if (!buf) return;
    A.side = blockIdx.x ? 1 : -1;
    slab_unpack_body(A, what, buf);
}

__global__ void k_slab_flags(Ctl *ctl, int hl, int hr) {
    ctl->halo_left = hl; ctl->halo_right = hr;
    ctl->halo_thick[0] = INFINITY; ctl->halo_thick[1] = INFINITY;
}

// ------------------------------------------------------------------------------------
// Conservation diagnostics over the alive particles (the checks the reference's author wished for,
// readme.md:23): sum m, sum m v, sum m v^2 / 2, sum m u.  Two stages with a fixed summation tree, so the
// numbers are reproducible run to run: CTA partials, then one CTA adds the partials in order.
constexpr int kDiagBlocks = 1024;
__global__ void __launch_bounds__(256) k_diag_partial(const double *m, const double *v, const double *u, int pad, int na,
                                                      double *partial) {
    double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < na; i += gridDim.x * blockDim.x) {
        const double mi = m[pad + i], vi = v[pad + i], ui = u[pad + i];
        s0 += mi; s1 = fma(mi, vi, s1); s2 = fma(0.5 * mi, vi * vi, s2); s3 = fma(mi, ui, s3);
    }
    __shared__ double sh[4][256];
    sh[0][threadIdx.x] = s0; sh[1][threadIdx.x] = s1; sh[2][threadIdx.x] = s2; sh[3][threadIdx.x] = s3;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o)
            for (int q = 0; q < 4; q++) sh[q][threadIdx.x] += sh[q][threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x < 4) partial[threadIdx.x * kDiagBlocks + blockIdx.x] = sh[threadIdx.x][0];
}
__global__ void __launch_bounds__(256) k_diag_final(const double *partial, int nblocks, double *out) {
    __shared__ double sh[4][256];
    for (int q = 0; q < 4; q++) {
        double s = 0;
        for (int b = threadIdx.x; b < nblocks; b += 256) s += partial[q * kDiagBlocks + b];
        sh[q][threadIdx.x] = s;
    }
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o)
            for (int q = 0; q < 4; q++) sh[q][threadIdx.x] += sh[q][threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x < 4) out[threadIdx.x] = sh[threadIdx.x][0];
}

// ------------------------------------------------------------------------------------
// Statistics of the most recent density pass over the alive particles, reduced on the device: Newton
// iteration histogram (12 bins, the last one open), neighbour count minimum / maximum / sum.
// out[0..11] histogram, out[12] sum, out[13] min, out[14] max (all unsigned long long, zeroed / preset by the caller).
__global__ void __launch_bounds__(256) k_count_stats(const int *cnt, int pad, int na, unsigned long long *out) {
    __shared__ unsigned hist[12];
    __shared__ unsigned long long ssum;
    __shared__ int smin, smax;
    if (threadIdx.x < 12) hist[threadIdx.x] = 0u;
    if (threadIdx.x == 0) { ssum = 0ull; smin = 0x7fffffff; smax = 0; }
    __syncthreads();
    unsigned long long sum = 0ull;
    int mn = 0x7fffffff, mx = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < na; i += gridDim.x * blockDim.x) {
        const int c = cnt[pad + i];
        const int nb = c & 0xffffff;
        atomicAdd(&hist[min((c >> 24) & 0xff, 11)], 1u);
        sum += (unsigned long long)nb; mn = min(mn, nb); mx = max(mx, nb);
    }
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_down_sync(0xffffffffu, sum, o);
    mn = __reduce_min_sync(0xffffffffu, mn);
    mx = __reduce_max_sync(0xffffffffu, mx);
    if ((threadIdx.x & 31) == 0) { atomicAdd(&ssum, sum); atomicMin(&smin, mn); atomicMax(&smax, mx); }
    __syncthreads();
    if (threadIdx.x < 12 && hist[threadIdx.x]) atomicAdd(out + threadIdx.x, (unsigned long long)hist[threadIdx.x]);
    if (threadIdx.x == 0) {
        atomicAdd(out + 12, ssum);
        atomicMin(out + 13, (unsigned long long)smin);
        atomicMax(out + 14, (unsigned long long)smax);
    }
}

// ------------------------------------------------------------------------------------
// A drift rarely breaks the x order by more than a few adjacent swaps.  One odd-even transposition
// step over the alive block: pairs (s, s+1) with s = pad + 2t + parity that are out of order swap every
// per-particle array that is live between the drift and the density pass.  A few steps of this repair
// such a step for the cost of reading x; the radix sort remains for what they cannot repair.
struct SwapArgs {
    double *x, *vh, *uh, *a, *du, *h, *rho, *P, *m;
    unsigned *idx;
    int pad, na;
};
__global__ void k_swap_inversions(SwapArgs A, int parity) {
    const int q = 2 * (blockIdx.x * blockDim.x + threadIdx.x) + parity;
    if (q + 1 >= A.na) return;
    const int s = A.pad + q;
    if (!(A.x[s] > A.x[s + 1])) return;
    double *arrs[9] = {A.x, A.vh, A.uh, A.a, A.du, A.h, A.rho, A.P, A.m};
#pragma unroll
    for (int k = 0; k < 9; k++) { const double t = arrs[k][s]; arrs[k][s] = arrs[k][s + 1]; arrs[k][s + 1] = t; }
    const unsigned ti = A.idx[s]; A.idx[s] = A.idx[s + 1]; A.idx[s + 1] = ti;
}

// ------------------------------------------------------------------------------------
// sort support
__global__ void k_iota(int *p, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = i;
}
__global__ void k_gather_f64(const double *in, double *out, const int *perm, int pad, int na) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < na) out[pad + i] = in[pad + perm[i]];
}
__global__ void k_gather_u32(const unsigned *in, unsigned *out, const int *perm, int pad, int na) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < na) out[pad + i] = in[pad + perm[i]];
}

// ------------------------------------------------------------------------------------
// t = 0 lattice (setup.cpp:145-185) and the adiabatic pre-pass's pressure / +-c_s
// velocities (setup.cpp:197-201).
struct LatticeArgs {
    double *x, *v, *u, *a, *du, *h, *rho, *P, *m;
    unsigned *idx;
    int *slot_id;
    int pad, na;
    long long i0;      // global index of this handle's first particle (x-slabs)
    double min_x, spacing, v0, u0, h0, mass;
};
__global__ void k_lattice(LatticeArgs A) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= A.na) return;
    const int s = A.pad + i;
    const double pos = __dadd_rn(A.min_x, __dmul_rn(A.spacing, (double)(A.i0 + i)));
    A.x[s] = pos;
    A.v[s] = (pos < 0) ? A.v0 : -A.v0;
    A.u[s] = A.u0;
    A.h[s] = A.h0;
    A.m[s] = A.mass;
    A.a[s] = 0; A.du[s] = 0; A.rho[s] = 0; A.P[s] = 0;
    A.idx[s] = (unsigned)(A.i0 + i);
    A.slot_id[i] = (int)(A.i0 + i);
}
// A DEVELOPED state of the same problem (synthetic; no counterpart in the reference, whose only start state is the
// lattice): the exact solution of the colliding adiabatic streams at time t, sampled with equal-mass particles.  Per
// half domain, from the wall inwards: gas at rest next to the wall (rho_w, c_w), a centred rarefaction fan
// (v + c = xi / t, v - 3 c = K for gamma = 5/3), the undisturbed stream (rho0, v0, c0), and the shocked gas at rest
// (rho1, P1) behind a shock moving outwards at speed S.  Particle i sits where the cumulative mass from the left wall is
// (i + 1/2) m; the right half is the mirror image.  h = h_factor m / rho, u = P / ((gamma - 1) rho).
struct DevelopedArgs {
    double *x, *v, *u, *a, *du, *h, *rho, *P, *m;
    unsigned *idx;
    int *slot_id;
    int pad, na;
    long long i0, n_global;
    double limit, mass, hfac, t;
    double rho0, P0, v0, c0, K, c_w, rho_w, P_w, S, rho1, P1;
    double M1, M2, M3, Mhalf;      // cumulative mass at the end of the wall region, the fan, the undisturbed stream; half the total
    double x_head;                 // where the fan ends (distance from the wall)
};
__global__ void k_developed(DevelopedArgs A) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= A.na) return;
    const int s = A.pad + i;
    const long long gi = A.i0 + i;
    const double mu = ((double)gi + 0.5) * A.mass;
    const bool right = mu > A.Mhalf;
    const double M = right ? 2.0 * A.Mhalf - mu : mu;       // mass between the particle and ITS wall
    double d, vel, rho, P;                                  // d = distance from that wall
    if (M <= A.M1) {
        d = M / A.rho_w; vel = 0.0; rho = A.rho_w; P = A.P_w;
    } else if (M <= A.M2) {
        // fan: mass(c) = M1 + rho0 t (c^4 - c_w^4) / c0^3
        const double c4 = fma((M - A.M1), (A.c0 * A.c0 * A.c0) / (A.rho0 * A.t), (A.c_w * A.c_w) * (A.c_w * A.c_w));
        const double c = sqrt(sqrt(c4));
        d = A.t * (4.0 * c + A.K);
        vel = (4.0 * c + A.K) - c;
        const double r = c / A.c0, r3 = r * r * r;
        rho = A.rho0 * r3; P = A.P0 * r3 * r * r;
    } else if (M <= A.M3) {
        d = A.x_head + (M - A.M2) / A.rho0; vel = A.v0; rho = A.rho0; P = A.P0;
    } else {
        d = (A.limit - A.S * A.t) + (M - A.M3) / A.rho1; vel = 0.0; rho = A.rho1; P = A.P1;
        d = fmin(d, A.limit);
    }
    A.x[s] = right ? A.limit - d : d - A.limit;
    A.v[s] = right ? -vel : vel;
    A.u[s] = P / ((kGamma - 1) * rho);
    A.h[s] = A.hfac * A.mass / rho;
    A.rho[s] = rho;
    A.P[s] = P;
    A.m[s] = A.mass;
    A.a[s] = 0; A.du[s] = 0;
    A.idx[s] = (unsigned)gi;
    A.slot_id[i] = (int)gi;
}
__global__ void k_prepass(const double *x, double *v, const double *u, const double *rho, double *P, int pad,
                          int na) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= na) return;
    const int s = pad + i;
    const double Pi = ((kGamma - 1) * u[s]) * rho[s];
    P[s] = Pi;
    const double c = sqrt(kGamma * Pi / rho[s]);
    v[s] = (x[s] < 0) ? c : -c;
}

__global__ void k_begin(Ctl *ctl, int clear_cand) {
    ctl->hmax_bits = 0ull;
    ctl->unsorted = 0;
    if (clear_cand) ctl->ncand = 0;
    ctl->nwork[0] = 0;
    ctl->nwork[1] = 0;
    ctl->nwork[2] = 0;
    ctl->nwork_peak = 0;
    ctl->tile_evals = 0;
    ctl->tiles_swept = 0;
    ctl->halo_changed = 0;
    ctl->sweep_flagged = 0;
    ctl->sweep_verify_fail = 0;
    ctl->sweep_moved_chunks = 0;
    for (int b = 0; b < kStatBins; b++) { ctl->force_pairs[b] = 0ull; ctl->density_scans[b] = 0ull; }
}

}  // namespace sphb
