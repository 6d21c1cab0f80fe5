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
constexpr int kHalo = 128;             // x staged in shared memory on each side of a tile

enum ErrCode {
    E_OK = 0, E_CAPACITY = 3, E_NEGATIVE_H = 4, E_LOW_DENSITY = 5, E_GHOST_ZERO_H = 6,
    E_ZERO_DERIVATIVE = 7, E_NOT_FINITE = 8, E_NO_BRACKET = 9, E_NOT_CONVERGED = 10,
    E_UNSORTED = 11
};

// Device control block (one per handle), read back by the host at sync points.
struct Ctl {
    int GL, GR;                    // ghosts currently in the left / right pad
    int ncand;                     // ghost candidates appended this step
    int unsorted;                  // order check failed
    int err_code, err_particle;    // first error wins
    unsigned long long hmax_bits;  // max over every h (old and new) in [lo,hi), as ordered bits
    unsigned long long fallbacks;  // bisection fallbacks taken
    int nchanged[2];               // particles whose kicked (v,u) changed, by sweep parity
    int sweeps_used;               // force sweeps the last step needed
    int pad_;
    unsigned long long force_pairs; // (i,j) pairs evaluated by the first force sweep (self included)
};

__device__ __forceinline__ void raise(Ctl *ctl, int code, int particle) {
    if (atomicCAS(&ctl->err_code, 0, code) == 0) ctl->err_particle = particle;
}

// ------------------------------------------------------------------------------------
// M5 quartic spline (kernel.cpp:14-64), branch-free: the three pieces are the clamped
// powers of (5/2-q), (3/2-q), (1/2-q).  Returns w(q); dw receives dw/dq.
__device__ __forceinline__ double m5(double q, double &dw) {
    double t1 = fmax(2.5 - q, 0.0), t2 = fmax(1.5 - q, 0.0), t3 = fmax(0.5 - q, 0.0);
    double s1 = t1 * t1, s2 = t2 * t2, s3 = t3 * t3;
    double w = fma(10.0, s3 * s3, fma(-5.0, s2 * s2, s1 * s1));
    double d = fma(-40.0, s3 * t3, fma(20.0, s2 * t2, -4.0 * (s1 * t1)));
    dw = kSigma * d;
    return kSigma * w;
}
__device__ __forceinline__ double m5_dw(double q) {
    double t1 = fmax(2.5 - q, 0.0), t2 = fmax(1.5 - q, 0.0), t3 = fmax(0.5 - q, 0.0);
    double d = fma(-40.0, (t3 * t3) * t3, fma(20.0, (t2 * t2) * t2, -4.0 * ((t1 * t1) * t1)));
    return kSigma * d;
}

// x window of a tile staged in shared memory, global fallback outside it.
struct XWin {
    const double *__restrict__ g;
    const double *sm;
    int base, n;
    __device__ __forceinline__ double operator()(int j) const {
        int k = j - base;
        return ((unsigned)k < (unsigned)n) ? sm[k] : __ldg(g + j);
    }
};

// ------------------------------------------------------------------------------------
// One evaluation of the two density-pass sums at trial smoothing length h:
//   sw = sum_j w(q_j)                      (calc_density, smoothing_length.cpp:80-91,
//                                           without the common factor m_i / h)
//   sd = sum_j m_j (w(q_j) + q_j w'(q_j))  (calc_density_dh, :53-62, without -1/h^2)
//   cnt = #{ j : |x_i - x_j| / h < 2.5 }   with the reference's IEEE division
// over the sorted window, self included.  The scan stops at the first q >= 2.5: the
// array is x-sorted, so every farther particle is outside the support as well.
template <bool UM>
__device__ __forceinline__ void density_sums(const XWin &X, const double *__restrict__ m, double m0,
                                             int s, int lo, int hi, double xi, double h,
                                             double &sw, double &sd, int &cnt) {
    const double ih = 1.0 / h;
    double dw0;
    double w0 = m5(0.0, dw0);
    double mi = UM ? m0 : __ldg(m + s);
    sw = w0;
    sd = mi * w0;
    cnt = 1;
    const double q_in = 2.5 * (1.0 - 0x1p-48), q_out = 2.5 * (1.0 + 0x1p-48);
#pragma unroll 1
    for (int side = 0; side < 2; side++) {
        int j = side ? s + 1 : s - 1;
        const int step = side ? 1 : -1;
#pragma unroll 1
        for (; side ? (j < hi) : (j >= lo); j += step) {
            double r = side ? (X(j) - xi) : (xi - X(j));
            double q = r * ih;
            if (q >= q_out) break;
            if (q > q_in) {                 // within rounding of the support edge: decide the
                if (r / h < kRadius) cnt++; // count exactly; the kernel value there is ~0
                else break;
                continue;
            }
            cnt++;
            double dw;
            double w = m5(q, dw);
            double mj = UM ? m0 : __ldg(m + j);
            sw += w;
            sd = fma(mj, fma(q, dw, w), sd);
        }
    }
}

struct DensityArgs {
    const double *x, *m, *h, *rho, *P;
    double *hn, *rhon, *omn, *oms, *Pn;
    int *cnt;            // (newton_iters << 24) | neighbour count
    Ctl *ctl;
    int pad, na;
    double limit, hfac, m0;
    int root_ghosts;     // 1: ghosts are root-found too (t = 0, setup.cpp:218-220)
    int root_alive;      // 0: omega-only pass (keeps h, rho; used by sphb_forces)
    int check_old;       // 1: the old (h, rho) state is also seen by force sums (step path)
};

// Fused density + variable-h root-find + omega, one thread per
// particle.  Newton follows rootfind_h (smoothing_length.cpp:191-259) and GSL's
// newton/test_delta step for step: same start value, same
// update, same stop test, same iteration cap, then the bisection fallback (:134-188).
// The last Newton evaluation is at the returned h, so its sums ARE the final density
// (calculators.cpp:45) and omega_new; the first evaluation is at the old h, so its
// derivative sum gives omega_stale = calc_omega(old h, old rho) at the new positions.
template <bool VARH, bool UM>
__global__ void __launch_bounds__(kTile) k_density(DensityArgs A) {
    __shared__ double sx[kTile + 2 * kHalo];
    Ctl *ctl = A.ctl;
    const int lo = A.pad - ctl->GL, hi = A.pad + A.na + ctl->GR;
    const int t0 = blockIdx.x * kTile;          // grid covers [0, pad + na + pad)
    if (t0 >= hi || t0 + kTile <= lo) return;
    const int wbase = max(t0 - kHalo, lo);
    const int wend = min(t0 + kTile + kHalo, hi);
    for (int k = threadIdx.x; k < wend - wbase; k += kTile) sx[k] = A.x[wbase + k];
    __syncthreads();
    XWin X{A.x, sx, wbase, wend - wbase};

    const int s = t0 + threadIdx.x;
    double hloc = 0.0;
    if (s >= lo && s < hi) {
        const bool is_pad = (s < A.pad) || (s >= A.pad + A.na);
        const bool do_root = VARH && (is_pad ? A.root_ghosts : A.root_alive);
        const double xi = X(s);
        const double mi = UM ? A.m0 : A.m[s];
        const double hold = A.h[s], rhoold = A.rho[s];
        const int npart = A.na + ctl->GL + ctl->GR;
        double hnew = hold, rhonew = rhoold, om_new = 1.0, om_stale = 1.0;
        int iters = 0, cnt = 0;
        double sw, sd;
        if (VARH) {
            double x0 = hold > kEps ? hold : A.hfac * (2 * A.limit / (npart - 1));
            density_sums<UM>(X, A.m, A.m0, s, lo, hi, xi, x0, sw, sd, cnt);
            double ih = 1.0 / x0;
            double f = mi * (sw * ih) - mi * A.hfac * ih;
            double dsum = -(sd * ih * ih);                    // sum_j m_j dW/dh
            double df = dsum + mi * A.hfac * ih * ih;
            om_stale = 1.0 - dsum * (-hold / rhoold);
            if (do_root) {
                double xcur = x0;
                bool cont = true;
#pragma unroll 1
                while (cont && iters < kMaxIterNR) {
                    iters++;
                    if (df == 0.0) { raise(ctl, E_ZERO_DERIVATIVE, s); break; }
                    double xn = xcur - f / df;
                    density_sums<UM>(X, A.m, A.m0, s, lo, hi, xi, xn, sw, sd, cnt);
                    ih = 1.0 / xn;
                    f = mi * (sw * ih) - mi * A.hfac * ih;
                    dsum = -(sd * ih * ih);
                    df = dsum + mi * A.hfac * ih * ih;
                    if (!isfinite(f) || !isfinite(df)) { raise(ctl, E_NOT_FINITE, s); break; }
                    double xprev = xcur;
                    xcur = xn;
                    cont = !(fabs(xcur - xprev) < kHEps * fabs(xcur) || xcur == xprev);
                }
                if (cont && ctl->err_code == 0) {
                    // bisection fallback on [1e-8, 2 limit], gsl_root_test_interval rel 1e-4
                    atomicAdd(&ctl->fallbacks, 1ull);
                    double xl = kEps, xh = 2 * A.limit, fl, fh;
                    int c2;
                    density_sums<UM>(X, A.m, A.m0, s, lo, hi, xi, xl, sw, sd, c2);
                    fl = mi * (sw / xl) - mi * A.hfac / xl;
                    density_sums<UM>(X, A.m, A.m0, s, lo, hi, xi, xh, sw, sd, c2);
                    fh = mi * (sw / xh) - mi * A.hfac / xh;
                    double root = 0.5 * (xl + xh);
                    if (!isfinite(fl) || !isfinite(fh)) raise(ctl, E_NOT_FINITE, s);
                    else if ((fl < 0.0 && fh < 0.0) || (fl > 0.0 && fh > 0.0)) raise(ctl, E_NO_BRACKET, s);
                    else {
                        int it = 0;
                        bool go = true;
#pragma unroll 1
                        while (go && it < kMaxIterBS) {
                            it++;
                            if (fl == 0.0) { root = xl; xh = xl; }
                            else if (fh == 0.0) { root = xh; xl = xh; }
                            else {
                                double mid = (xl + xh) / 2.0;
                                density_sums<UM>(X, A.m, A.m0, s, lo, hi, xi, mid, sw, sd, c2);
                                double fm = mi * (sw / mid) - mi * A.hfac / mid;
                                if (!isfinite(fm)) { raise(ctl, E_NOT_FINITE, s); break; }
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
                    xcur = root;
                    density_sums<UM>(X, A.m, A.m0, s, lo, hi, xi, xcur, sw, sd, cnt);
                    ih = 1.0 / xcur;
                    dsum = -(sd * ih * ih);
                }
                if (xcur < 0) raise(ctl, E_NEGATIVE_H, s);
                hnew = xcur;
                rhonew = mi * (sw * ih);
                om_new = 1.0 - dsum * (-hnew / rhonew);
            } else {
                om_new = om_stale;   // particle keeps its stored h, rho: omega for those
            }
        } else {
            density_sums<UM>(X, A.m, A.m0, s, lo, hi, xi, hold, sw, sd, cnt);
            if (is_pad ? A.root_ghosts : A.root_alive) rhonew = mi * (sw / hold);
        }
        A.hn[s] = hnew;
        A.rhon[s] = rhonew;
        A.omn[s] = om_new;
        A.oms[s] = om_stale;
        if (A.root_alive || A.root_ghosts) A.cnt[s] = (iters << 24) | (cnt & 0xffffff);
        if (is_pad && A.Pn != A.P) A.Pn[s] = A.P[s];
        // ensure_nonzero_density (calculators.cpp:174-186) guards every particle a force sum
        // touches, in whichever state (new, or old when check_old) it will be seen
        if (rhonew < kEps || (A.check_old && rhoold < kEps)) raise(ctl, E_LOW_DENSITY, s);
        hloc = fmax(hnew, hold);
    }
    // block max of h -> global upper bound for the force pass window
    __shared__ double smax[kTile / 32];
    for (int o = 16; o > 0; o >>= 1) hloc = fmax(hloc, __shfl_xor_sync(0xffffffffu, hloc, o));
    if ((threadIdx.x & 31) == 0) smax[threadIdx.x >> 5] = hloc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double mx = smax[0];
        for (int k = 1; k < kTile / 32; k++) mx = fmax(mx, smax[k]);
        atomicMax(&ctl->hmax_bits, (unsigned long long)__double_as_longlong(mx));
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
    const double *hA, *rA, *oA, *vA, *uA;
    const double *hB, *rB, *oB, *vB, *uB;
};

enum { SEL_INDEX = 0, SEL_ALIVE = 1, SEL_ALL_A = 2 };

template <bool VOL>
__device__ __forceinline__ double ld(const double *p) {
    return VOL ? __ldcg(p) : __ldg(p);
}

template <int SEL, bool ADIA, bool UM, bool VOL>
__device__ __forceinline__ void force_eval(const ForceArrays &F, double m0, int s, int lo, int hi,
                                           double Hb, double xi, unsigned idxi, double hi_, double rhoi,
                                           double omi, double vi, double ui, double Pstale,
                                           double &acc, double &dudt, double &Pi_out, int &jmin, int &jmax, int &npairs) {
    const double Pi = ADIA ? ((kGamma - 1) * ui) * rhoi : rhoi;
    const double c_ac = ADIA ? sqrt(kGamma * Pstale / rhoi) : 1.0;   // stale stored pressure (F7)
    const double c_ec = ADIA ? sqrt(kGamma * Pi / rhoi) : 1.0;       // fresh pressure
    const double PrRho_i = Pi / (rhoi * rhoi) / omi;
    const double ihi = 1.0 / hi_;
    const double ihi2 = ihi * ihi;
    const double reach = kRadius * Hb;
    double a_sum = 0.0, e_sum = 0.0;
    jmin = s; jmax = s;
    Pi_out = Pi;
    npairs = 1;   // self: its energy term is evaluated by the reference too (and is exactly 0)
#pragma unroll 1
    for (int side = 0; side < 2; side++) {
        int j = side ? s + 1 : s - 1;
        const int step = side ? 1 : -1;
#pragma unroll 1
        for (; side ? (j < hi) : (j >= lo); j += step) {
            const double xj = __ldg(F.x + j);
            const double r = xi - xj;
            const double ar = fabs(r);
            if (ar >= reach) break;
            if (side) jmax = j; else jmin = j;
            const unsigned idxj = __ldg(F.idx + j);
            bool useA;
            if (SEL == SEL_INDEX) useA = idxj < idxi;
            else if (SEL == SEL_ALIVE) useA = !(idxj & kGhostBit);
            else useA = true;
            const double hj = useA ? ld<false>(F.hA + j) : __ldg(F.hB + j);
            if (ar >= kRadius * fmax(hi_, hj)) continue;   // all three gradients vanish
            npairs++;
            const double rhoj = useA ? __ldg(F.rA + j) : __ldg(F.rB + j);
            const double omj = useA ? __ldg(F.oA + j) : __ldg(F.oB + j);
            const double vj = useA ? ld<VOL>(F.vA + j) : __ldg(F.vB + j);
            const double uj = useA ? ld<VOL>(F.uA + j) : __ldg(F.uB + j);
            const double mj = UM ? m0 : __ldg(F.m + j);
            // grad_W (calculators.cpp:17-27) with h_i, h_j and the mean h; r == 0 counts as "<= 0"
            const double sg = (r > 0) ? 1.0 : -1.0;
            const double hij = (hi_ + hj) / 2;
            const double v_ij = vi - vj;
            const double dot = v_ij * r;
            const double rho_sum = rhoi + rhoj;
            const double D = fma(r, r, 0.01 * (hij * hij));
            // one reciprocal serves 1/h_ij, 1/D and 1/rho_sum
            const double R = 1.0 / (hij * D * rho_sum);
            const double ihij = R * (D * rho_sum);
            const double iD = R * (hij * rho_sum);
            const double irs = R * (hij * D);
            const double ihj = 1.0 / hj;
            const double gWi = sg * (m5_dw(ar * ihi) * ihi2);
            const double gWj = sg * (m5_dw(ar * ihj) * (ihj * ihj));
            const double gWij = sg * (m5_dw(ar * ihij) * (ihij * ihij));
            const double Pj = ADIA ? ((kGamma - 1) * uj) * rhoj : rhoj;
            const double PrRho_j = Pj / (rhoj * rhoj) / omj;
            // artificial viscosity (calculators.cpp:148-172): alpha = 1, beta = 2
            double visc_ac = 0.0, visc_ec = 0.0;
            if (!(dot > 0)) {
                const double mu = (hij * dot) * iD;
                const double two_mu2 = 2.0 * (mu * mu);
                const double irho = 2.0 * irs;             // 1 / ((rho_i + rho_j) / 2)
                visc_ac = fma(-c_ac, mu, two_mu2) * irho;
                visc_ec = fma(-c_ec, mu, two_mu2) * irho;
            }
            if (idxj != idxi)
                a_sum -= mj * fma(gWij, visc_ac, fma(gWj, PrRho_j, gWi * PrRho_i));
            const double mv = mj * v_ij;
            e_sum = fma(PrRho_i * mv, gWij, e_sum);
            e_sum = fma((0.5 * mv) * visc_ec, gWij, e_sum);
        }
    }
    acc = a_sum;
    dudt = e_sum;
}

struct SweepArgs {
    ForceArrays F;
    const double *Pstale;   // stored pressure of the previous force pass
    double *Pout;           // new stored pressure
    double *a, *du;         // outputs
    double *vn, *un;        // fully-kicked outputs (iterate)
    const double *vh, *uh;  // half-kicked base state of i itself
    const double *hO, *rO, *oO;   // own h, rho, omega of i
    int *dirty;             // per-tile stamp: epoch of the last sweep that changed something in range
    Ctl *ctl;
    int pad, na;
    double m0, half_dt;
    int epoch;              // stamp of this sweep
    int kick;               // 0: t = 0 semantics (no kick, calculators only)
};

__device__ __forceinline__ void mark_dirty(int *dirty, int pad, int na, int jmin, int jmax, int epoch) {
    int ntiles = (na + kTile - 1) / kTile;
    int t0 = max((jmin - pad) / kTile, 0), t1 = min((jmax - pad) / kTile, ntiles - 1);
    if (jmin < pad) t0 = 0;
    for (int t = t0; t <= t1; t++) dirty[t] = epoch;
}

// First force sweep of a step over every alive particle (also the only one for the
// jacobi order and for the kick-less t = 0 pass).
template <int SEL, bool ADIA, bool UM>
__global__ void __launch_bounds__(kTile) k_force_first(SweepArgs A) {
    Ctl *ctl = A.ctl;
    const int lo = A.pad - ctl->GL, hi = A.pad + A.na + ctl->GR;
    const int s = A.pad + blockIdx.x * kTile + threadIdx.x;
    bool changed = false;
    int npairs = 0;
    if (s < A.pad + A.na) {
        const double Hb = __longlong_as_double((long long)ctl->hmax_bits);
        double acc, dudt, Pi;
        int jmin, jmax;
        const double vi = A.vh[s], ui = A.uh[s];
        force_eval<SEL, ADIA, UM, false>(A.F, A.m0, s, lo, hi, Hb, A.F.x[s], A.F.idx[s], A.hO[s], A.rO[s],
                                         A.oO[s], vi, ui, A.Pstale[s], acc, dudt, Pi, jmin, jmax, npairs);
        A.a[s] = acc;
        A.du[s] = dudt;
        A.Pout[s] = Pi;
        if (A.kick) {
            // second half-kick, sph_simulation.cpp:105-106, operation by operation
            const double vf = __dadd_rn(vi, __dmul_rn(acc, A.half_dt));
            const double uf = __dadd_rn(ui, __dmul_rn(dudt, A.half_dt));
            A.vn[s] = vf;
            A.un[s] = uf;
            if (SEL == SEL_INDEX) {
                // later-index neighbours used (vh, uh) as a guess for this particle
                changed = (__double_as_longlong(vf) != __double_as_longlong(vi)) ||
                          (__double_as_longlong(uf) != __double_as_longlong(ui));
                if (changed) mark_dirty(A.dirty, A.pad, A.na, jmin, jmax, A.epoch);
            }
        }
    }
    if (SEL == SEL_INDEX && A.kick) {
        if (__syncthreads_or(changed) && threadIdx.x == 0) atomicAdd(&ctl->nchanged[A.epoch & 1], 1);
    }
    // pair statistics for the roofline accounting (one atomic per warp)
    for (int o = 16; o > 0; o >>= 1) npairs += __shfl_xor_sync(0xffffffffu, npairs, o);
    if ((threadIdx.x & 31) == 0 && npairs) atomicAdd(&ctl->force_pairs, (unsigned long long)npairs);
}

// Refinement of the ordered sweep: ONE persistent cooperative launch that repeats
// "recompute every tile in range of a particle whose kicked (v,u) changed in the
// previous sweep" until a whole sweep changes nothing.  The dependency of particle i on
// the fully-kicked (v,u) of its lower-index neighbours is strictly lower triangular, so
// this fixed point exists, is unique, is reached after finitely many sweeps and IS the
// result of the reference's sequential in-place sweep.
// Updates are made in place (chaotic relaxation); a sweep that changes nothing proves
// that every particle was last computed from the final values.
template <bool ADIA, bool UM>
__global__ void __launch_bounds__(kTile) k_force_refine(SweepArgs A, int max_sweeps) {
    cg::grid_group grid = cg::this_grid();
    Ctl *ctl = A.ctl;
    const int lo = A.pad - ctl->GL, hi = A.pad + A.na + ctl->GR;
    const int ntiles = (A.na + kTile - 1) / kTile;
    const double Hb = __longlong_as_double((long long)ctl->hmax_bits);
    int epoch = A.epoch;          // stamp of the first sweep (k_force_first)
    int used = 1;
    while (true) {
        const int prev_changed = *((volatile int *)&ctl->nchanged[epoch & 1]);
        if (prev_changed == 0) break;
        if (used >= max_sweeps) {
            if (blockIdx.x == 0 && threadIdx.x == 0) raise(ctl, E_NOT_CONVERGED, -1);
            break;
        }
        epoch++;
        used++;
        if (blockIdx.x == 0 && threadIdx.x == 0) ctl->nchanged[epoch & 1] = 0;
        grid.sync();
        for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
            if (__ldcg(A.dirty + t) < epoch - 1) continue;
            const int s = A.pad + t * kTile + threadIdx.x;
            bool changed = false;
            if (s < A.pad + A.na) {
                double acc, dudt, Pi;
                int jmin, jmax, npairs;
                const double vi = A.vh[s], ui = A.uh[s];
                force_eval<SEL_INDEX, ADIA, UM, true>(A.F, A.m0, s, lo, hi, Hb, A.F.x[s], A.F.idx[s], A.hO[s],
                                                      A.rO[s], A.oO[s], vi, ui, A.Pstale[s], acc, dudt, Pi,
                                                      jmin, jmax, npairs);
                const double vf = __dadd_rn(vi, __dmul_rn(acc, A.half_dt));
                const double uf = __dadd_rn(ui, __dmul_rn(dudt, A.half_dt));
                const double vo = __ldcg(A.vn + s), uo = __ldcg(A.un + s);
                changed = (__double_as_longlong(vf) != __double_as_longlong(vo)) ||
                          (__double_as_longlong(uf) != __double_as_longlong(uo));
                if (changed) {
                    A.a[s] = acc;
                    A.du[s] = dudt;
                    __stcg(A.vn + s, vf);
                    __stcg(A.un + s, uf);
                    mark_dirty(A.dirty, A.pad, A.na, jmin, jmax, epoch);
                }
            }
            if (__syncthreads_or(changed) && threadIdx.x == 0) atomicAdd(&ctl->nchanged[epoch & 1], 1);
        }
        grid.sync();
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) ctl->sweeps_used = used;
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
        G.a[dest] = G.ga[c];
        G.du[dest] = G.gdu[c];
        G.h[dest] = G.gh[c];
        G.rho[dest] = G.grho[c];
        G.P[dest] = G.gP[c];
        G.m[dest] = G.gm[c];
        G.idx[dest] = kGhostBit | kc;
        G.gorder[rank_k] = dest;
    }
    if (threadIdx.x == 0) { ctl->GL = GL; ctl->GR = GR; }
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
};

__global__ void k_unpack(PackArgs A) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= A.na + A.ng) return;
    const AosRec r = A.aos[i];
    A.slot_id[i] = r.id;
    if (i < A.na) {
        const int s = A.pad + i;
        A.x[s] = r.pos; A.v[s] = r.vel; A.u[s] = r.u; A.a[s] = r.acc; A.du[s] = r.du_dt;
        A.h[s] = r.h; A.rho[s] = r.density; A.P[s] = r.pressure; A.m[s] = r.mass;
        A.idx[s] = (unsigned)i;
    } else {
        const int k = i - A.na;
        A.gx[k] = r.pos; A.gv[k] = r.vel; A.gu[k] = r.u; A.ga[k] = r.acc; A.gdu[k] = r.du_dt;
        A.gh[k] = r.h; A.grho[k] = r.density; A.gP[k] = r.pressure; A.gm[k] = r.mass;
        A.gkey[k] = (unsigned)k;
    }
}

__global__ void k_pack(PackArgs A) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int ng = A.ctl->GL + A.ctl->GR;
    if (t >= A.na + ng) return;
    int s, out;
    const double *vv, *uu;
    int type;
    if (t < A.na) {
        s = A.pad + t;
        out = (int)A.idx[s];
        vv = A.v; uu = A.u; type = 0;
    } else {
        s = A.gorder[t - A.na];
        out = t;
        vv = A.vg; uu = A.ug; type = 1;
    }
    AosRec r;
    r.id = A.slot_id[out]; r.pad0 = 0;
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
                           int pad, int na) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int ng = ctl->GL + ctl->GR;
    if (t >= na + ng) return;
    if (t < na) {
        const int s = pad + t;
        out[idx[s]] = src[s];
    } else {
        out[t] = src[gorder[t - na]];
    }
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
    double min_x, spacing, v0, u0, h0, mass;
};
__global__ void k_lattice(LatticeArgs A) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= A.na) return;
    const int s = A.pad + i;
    const double pos = __dadd_rn(A.min_x, __dmul_rn(A.spacing, (double)i));
    A.x[s] = pos;
    A.v[s] = (pos < 0) ? A.v0 : -A.v0;
    A.u[s] = A.u0;
    A.h[s] = A.h0;
    A.m[s] = A.mass;
    A.a[s] = 0; A.du[s] = 0; A.rho[s] = 0; A.P[s] = 0;
    A.idx[s] = (unsigned)i;
    A.slot_id[i] = i;
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
    ctl->nchanged[0] = 0;
    ctl->nchanged[1] = 0;
    ctl->force_pairs = 0ull;
}

}  // namespace sphb
