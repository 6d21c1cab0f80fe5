/*
 * oracle/sph_oracle.c — TEST INFRASTRUCTURE, not product code.
 *
 * Plain-C O(N·k) restatement of the reference's per-timestep particle loop.
 * The reference loops over ALL particles in every sum and relies on its kernel
 * returning 0 beyond q = 2.5 (reference sph/kernel.cpp:34-36); here the
 * particles are x-sorted once per step and every sum runs over the bounded
 * window of candidates that can be non-zero.  Skipping an exactly-zero term
 * does not change an IEEE sum, so in index_order_sums mode (terms accumulated
 * in ascending array index, like the reference) the results are bit-identical
 * to the reference's own strict build; tests/test_oracle_golden.py checks that
 * against fixtures those builds produced (tests/golden/make_golden.py).
 *
 * Reference map (all paths relative to /root/reference/):
 *   orc_w / orc_dw            sph/kernel.cpp:14-38, 40-64
 *   density_sum               sph/smoothing_length.cpp:80-91   (calc_density)
 *   density_dh_sum            sph/smoothing_length.cpp:28-50,53-62 (calc_dW_dh, calc_density_dh)
 *   omega_of                  sph/smoothing_length.cpp:64-77   (calc_omega)
 *   eval_f_df                 sph/smoothing_length.cpp:94-130  (smoothing_f/df/fdf)
 *   rootfind_h                sph/smoothing_length.cpp:191-259 + GSL newton
 *   rootfind_h_fallback       sph/smoothing_length.cpp:134-188 + GSL bisection
 *   density_one               sph/calculators.cpp:33-46 (variable h), :51-54 (fixed h)
 *   grad_w                    sph/calculators.cpp:17-27
 *   viscosity                 sph/calculators.cpp:148-172; alpha, beta, eta: calculators.hpp:62-64
 *   acceleration_one          sph/calculators.cpp:63-129, EOS :131-137, sound speed :139-146
 *   energy_one                sph/calculators.cpp:192-213
 *   orc_ghosts                sph/ghost_particles.cpp:19-115
 *   orc_kick_drift, orc_step  sph/sph_simulation.cpp:55-108
 *   orc_init_particles        sph/setup.cpp:143-228
 *   constants                 sph/define.hpp:25,29,32,39,41,43; sph/kernel.hpp:12
 *
 * Build: gcc -std=c11 -O2 -ffp-contract=off (oracle/Makefile).
 */
#include "sph_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define CALC_EPSILON 1e-8
#define GAMMA (5.0 / 3.0)
#define H_MAX_ITER_NR 10
#define H_MAX_ITER_BS 1000
#define H_EPSILON 1e-4
#define KERNEL_RADIUS 2.5
#define AV_ALPHA 1.0
#define AV_BETA 2.0
#define AV_ETA 0.01
#define WIN_MARGIN (1.0 + 1e-12)

struct orc_sim {
    /* Config (reference sph/basictypes.hpp:21-31) */
    int n_part;      /* alive + ghost, like Config::n_part after ghost set-up */
    int n_ghost;
    double cfg_mass, limit, v_0, h_factor, t_i;
    int pressure_calc;
    /* compile-time switches of the reference turned into run-time options */
    int variable_h;
    double constant_h;
    int order;
    int index_order_sums;
    /* particle arrays, array-index order: alive block then ghost block */
    int cap;
    double *mass, *pos, *vel, *acc, *h, *du_dt, *u, *density, *pressure;
    /* x-sorted view, rebuilt when positions or the ghost set change */
    int *perm;       /* sorted rank -> array index */
    int *rank;       /* array index -> sorted rank */
    double *xs;      /* positions in sorted order */
    int *alive_perm; /* previous alive ordering (adaptive re-sort) */
    int alive_perm_n;
    double hmax;     /* upper bound of every h currently stored */
    /* caches / outputs */
    double *omega;   /* omega of each particle for its current-in-array h, density */
    int *newton_iters;
    int *nbr_count;
    long fallback_count;
    double force_nbr_sum;
    long force_nbr_n;
    int *scratch;    /* candidate list */
    int scratch_cap;
    int err;
    char errmsg[256];
};

/* ------------------------------------------------------------------ utils */

static int fail(orc_sim *s, int code, const char *msg, int id, double v) {
    s->err = code;
    snprintf(s->errmsg, sizeof s->errmsg, "%s (particle %d, value %.17g)", msg, id, v);
    return code;
}

static int ensure_cap(orc_sim *s, int need) {
    if (need <= s->cap) return 0;
    int ncap = s->cap ? s->cap : 16;
    while (ncap < need) ncap += ncap / 2 + 16;
    double **arrs[] = {&s->mass, &s->pos, &s->vel, &s->acc, &s->h, &s->du_dt, &s->u,
                       &s->density, &s->pressure, &s->xs, &s->omega};
    for (size_t k = 0; k < sizeof arrs / sizeof arrs[0]; k++) {
        double *p = (double *)realloc(*arrs[k], sizeof(double) * (size_t)ncap);
        if (!p) return ORC_ERR_ALLOC;
        for (int i = s->cap; i < ncap; i++) p[i] = 0.0;
        *arrs[k] = p;
    }
    int **iarrs[] = {&s->perm, &s->rank, &s->alive_perm, &s->newton_iters, &s->nbr_count};
    for (size_t k = 0; k < sizeof iarrs / sizeof iarrs[0]; k++) {
        int *p = (int *)realloc(*iarrs[k], sizeof(int) * (size_t)ncap);
        if (!p) return ORC_ERR_ALLOC;
        for (int i = s->cap; i < ncap; i++) p[i] = 0;
        *iarrs[k] = p;
    }
    s->cap = ncap;
    return 0;
}

static int n_alive(const orc_sim *s) { return s->n_part - s->n_ghost; }

/* -------------------------------------------------- kernel.cpp:14-38,40-64 */

double orc_w(double q) {
    double w;
    double sigma = 1. / 24.;
    if (q < 0.5) {
        w = pow(5. / 2. - q, 4) - 5 * pow(3. / 2. - q, 4) + 10 * pow(1. / 2. - q, 4);
        return sigma * w;
    } else if (q >= 0.5 && q < 3. / 2.) {
        w = pow(5. / 2. - q, 4) - 5 * pow(3. / 2. - q, 4);
        return sigma * w;
    } else if (q >= 3. / 2. && q < 5. / 2.) {
        w = pow(5. / 2. - q, 4);
        return sigma * w;
    }
    return 0;
}

double orc_dw(double q) {
    double w;
    double sigma = 1. / 24.;
    if (q < 0.5) {
        w = -4 * pow(5. / 2. - q, 3) + 20 * pow(3. / 2. - q, 3) - 40 * pow(1. / 2. - q, 3);
        return sigma * w;
    } else if (q >= 0.5 && q < 3. / 2.) {
        w = -4 * pow(5. / 2. - q, 3) + 20 * pow(3. / 2. - q, 3);
        return sigma * w;
    } else if (q >= 3. / 2. && q < 5. / 2.) {
        w = -4 * pow(5. / 2. - q, 3);
        return sigma * w;
    }
    return 0;
}

/* ------------------------------------------------------------ sorted view */

static void sort_small_by_pos(const double *pos, int *idx, int n) {
    for (int a = 1; a < n; a++) {
        int v = idx[a];
        double x = pos[v];
        int b = a - 1;
        while (b >= 0 && (pos[idx[b]] > x || (pos[idx[b]] == x && idx[b] > v))) {
            idx[b + 1] = idx[b];
            b--;
        }
        idx[b + 1] = v;
    }
}

/* Rebuild perm/rank/xs for the current positions.  Alive particles move a
 * fraction of a spacing per step, so their previous order is re-used and
 * repaired by insertion; the handful of ghosts is sorted and merged in. */
static void rebuild_sorted(orc_sim *s) {
    int na = n_alive(s), ng = s->n_ghost, n = s->n_part;
    if (s->alive_perm_n != na) {
        for (int i = 0; i < na; i++) s->alive_perm[i] = i;
        s->alive_perm_n = na;
    }
    sort_small_by_pos(s->pos, s->alive_perm, na); /* O(N + inversions) */
    int *g = s->rank; /* borrow as temp for the ghost order */
    for (int k = 0; k < ng; k++) g[k] = na + k;
    sort_small_by_pos(s->pos, g, ng);
    int a = 0, b = 0, r = 0;
    /* copy ghost order out before rank[] is overwritten */
    int *gtmp = (int *)malloc(sizeof(int) * (size_t)(ng > 0 ? ng : 1));
    memcpy(gtmp, g, sizeof(int) * (size_t)ng);
    while (a < na || b < ng) {
        int take_alive;
        if (a >= na) take_alive = 0;
        else if (b >= ng) take_alive = 1;
        else take_alive = s->pos[s->alive_perm[a]] <= s->pos[gtmp[b]];
        s->perm[r++] = take_alive ? s->alive_perm[a++] : gtmp[b++];
    }
    free(gtmp);
    double hm = 0.0;
    for (r = 0; r < n; r++) {
        int i = s->perm[r];
        s->rank[i] = r;
        s->xs[r] = s->pos[i];
        if (s->h[i] > hm) hm = s->h[i];
    }
    s->hmax = hm;
}

/* Candidate array indices whose |x - x_i| can be below `radius`; always
 * contains i itself.  Ascending array index when index_order_sums. */
static int gather(orc_sim *s, int i, double radius, int **out) {
    int n = s->n_part;
    double xi = s->pos[i];
    double reach = radius * WIN_MARGIN;
    int r0 = s->rank[i];
    int lo = r0, hi = r0;
    while (lo > 0 && xi - s->xs[lo - 1] <= reach) lo--;
    while (hi + 1 < n && s->xs[hi + 1] - xi <= reach) hi++;
    int cnt = hi - lo + 1;
    if (cnt > s->scratch_cap) {
        int nc = cnt * 2 + 64;
        int *p = (int *)realloc(s->scratch, sizeof(int) * (size_t)nc);
        if (!p) return -1;
        s->scratch = p;
        s->scratch_cap = nc;
    }
    for (int r = lo; r <= hi; r++) s->scratch[r - lo] = s->perm[r];
    if (s->index_order_sums) {
        int *a = s->scratch;
        for (int p = 1; p < cnt; p++) {
            int v = a[p], b = p - 1;
            while (b >= 0 && a[b] > v) { a[b + 1] = a[b]; b--; }
            a[b + 1] = v;
        }
    }
    *out = s->scratch;
    return cnt;
}

/* ------------------------------------- smoothing_length.cpp:80-91, 28-62 */

/* calc_density: rho(h) = sum_j m_i * (w(q)/h), self and ghosts included, the
 * weight is the i-particle's OWN mass. */
static double density_sum(orc_sim *s, int i, double h, int *count_out) {
    int *c;
    int n = gather(s, i, KERNEL_RADIUS * h, &c);
    double d_sum = 0;
    int cnt = 0;
    double mi = s->mass[i], xi = s->pos[i];
    for (int k = 0; k < n; k++) {
        double q = fabs(xi - s->pos[c[k]]) / h;
        double w = orc_w(q);
        if (q < KERNEL_RADIUS) cnt++;
        d_sum += mi * (w / h);
    }
    if (count_out) *count_out = cnt;
    return d_sum;
}

/* calc_density_dh: sum_j m_j * dW/dh, dW/dh = (w + q w') / (-(h^2)) */
static double density_dh_sum(orc_sim *s, int i, double h) {
    int *c;
    int n = gather(s, i, KERNEL_RADIUS * h, &c);
    double d_sum = 0;
    double xi = s->pos[i];
    for (int k = 0; k < n; k++) {
        int j = c[k];
        double r_ij = xi - s->pos[j];
        double q = fabs(r_ij) / h;
        double dW_dh = (orc_w(q) + q * orc_dw(q)) / (-pow(h, 2));
        d_sum += s->mass[j] * dW_dh;
    }
    return d_sum;
}

/* calc_omega (smoothing_length.cpp:64-77) with the particle's stored h, rho */
static double omega_of(orc_sim *s, int i) {
    if (!s->variable_h) return 1;
    double o_sum = density_dh_sum(s, i, s->h[i]);
    double dh_drho = -s->h[i] / s->density[i];
    o_sum *= dh_drho;
    return 1 - o_sum;
}

/* ----------------------------------- smoothing_length.cpp:94-130 (f, f') */

static double eval_f(orc_sim *s, int i, double x) {
    double density_sum_v = density_sum(s, i, x, NULL);
    double density_exp = s->mass[i] * s->h_factor / x;
    return density_sum_v - density_exp;
}

static double eval_df(orc_sim *s, int i, double x) {
    double drho_dh_sum = density_dh_sum(s, i, x);
    double drho_dh_exp = -s->mass[i] * s->h_factor / pow(x, 2);
    return drho_dh_sum - drho_dh_exp;
}

/* smoothing_length.cpp:134-188 + GSL bisection semantics */
static int rootfind_h_fallback(orc_sim *s, int i, double *root_out) {
    double x_lo = CALC_EPSILON, x_hi = 2 * s->limit;
    double x = 0.5 * (x_lo + x_hi);
    double f_lo = eval_f(s, i, x_lo), f_hi = eval_f(s, i, x_hi);
    if (!isfinite(f_lo) || !isfinite(f_hi))
        return fail(s, ORC_ERR_NOT_FINITE, "bisection: function value is not finite", i, x_lo);
    if ((f_lo < 0.0 && f_hi < 0.0) || (f_lo > 0.0 && f_hi > 0.0))
        return fail(s, ORC_ERR_NO_BRACKET, "bisection: endpoints do not straddle y=0", i, f_lo);
    int iter = 0, cont;
    s->fallback_count++;
    do {
        iter++;
        if (f_lo == 0.0) {
            x = x_lo; x_hi = x_lo;
        } else if (f_hi == 0.0) {
            x = x_hi; x_lo = x_hi;
        } else {
            double mid = (x_lo + x_hi) / 2.0;
            double f_mid = eval_f(s, i, mid);
            if (!isfinite(f_mid))
                return fail(s, ORC_ERR_NOT_FINITE, "bisection: function value is not finite", i, mid);
            if (f_mid == 0.0) {
                x = mid; x_lo = mid; x_hi = mid;
            } else if ((f_lo > 0.0 && f_mid < 0.0) || (f_lo < 0.0 && f_mid > 0.0)) {
                x = 0.5 * (x_lo + mid); x_hi = mid; f_hi = f_mid;
            } else {
                x = 0.5 * (mid + x_hi); x_lo = mid; f_lo = f_mid;
            }
        }
        /* gsl_root_test_interval(lo, hi, 0, H_EPSILON) */
        double alo = fabs(x_lo), ahi = fabs(x_hi), smallest = 0.0;
        if ((x_lo > 0.0 && x_hi > 0.0) || (x_lo < 0.0 && x_hi < 0.0)) smallest = alo < ahi ? alo : ahi;
        cont = !(fabs(x_hi - x_lo) < H_EPSILON * smallest);
    } while (cont && iter < H_MAX_ITER_BS);
    *root_out = x;
    return 0;
}

/* smoothing_length.cpp:191-259 + GSL newton semantics */
static int rootfind_h(orc_sim *s, int i, double *root_out, int *iters_out) {
    double x0, x;
    if (s->h[i] > CALC_EPSILON) {
        x0 = s->h[i];
        x = s->h[i];
    } else {
        double mean_p_spacing = 2 * s->limit / (s->n_part - 1);
        x0 = s->h_factor * mean_p_spacing;
        x = s->h_factor * mean_p_spacing;
    }
    double f = eval_f(s, i, x), df = eval_df(s, i, x); /* fdfsolver_set */
    int iter = 0, cont;
    do {
        iter++;
        if (df == 0.0) return fail(s, ORC_ERR_ZERO_DERIVATIVE, "newton: derivative is zero", i, x);
        double next = x - (f / df);
        f = eval_f(s, i, next);
        df = eval_df(s, i, next);
        if (!isfinite(f) || !isfinite(df))
            return fail(s, ORC_ERR_NOT_FINITE, "newton: value not finite", i, next);
        x0 = x;
        x = next;
        /* gsl_root_test_delta(x, x0, 0, H_EPSILON) */
        cont = !(fabs(x - x0) < H_EPSILON * fabs(x) || x == x0);
    } while (cont && iter < H_MAX_ITER_NR);
    *iters_out = iter;
    if (cont) {
        int rc = rootfind_h_fallback(s, i, &x);
        if (rc) return rc;
    }
    *root_out = x;
    return 0;
}

/* DensityCalculator::operator() — calculators.cpp:33-46 / :51-54 */
static int density_one(orc_sim *s, int i) {
    if (s->variable_h) {
        double h = 0;
        int iters = 0;
        int rc = rootfind_h(s, i, &h, &iters);
        if (rc) return rc;
        s->newton_iters[i] = iters;
        if (h < 0) return fail(s, ORC_ERR_NEGATIVE_H, "root-finding returned negative smoothing length", i, h);
        s->h[i] = h;
        if (h > s->hmax) s->hmax = h;
    } else {
        s->newton_iters[i] = 0;
    }
    s->density[i] = density_sum(s, i, s->h[i], &s->nbr_count[i]);
    return 0;
}

/* --------------------------------------------- calculators.cpp:17-27 etc. */

static double grad_w(double r_ij, double h) {
    double r_ij_unit = (r_ij > 0) ? 1 : -1;
    double q = fabs(r_ij) / h;
    return (orc_dw(q) / pow(h, 2)) * r_ij_unit;
}

/* artificial_viscosity — calculators.cpp:148-172 */
static double viscosity(double v_i, double v_j, double rho_i, double rho_j, double r_ij, double h,
                        double c_s) {
    double v_ij = v_i - v_j;
    double dot = v_ij * r_ij;
    if (dot > 0) return 0;
    double eta_sq = AV_ETA * pow(h, 2);
    double mu_ij = (h * dot) / (pow(r_ij, 2) + eta_sq);
    double rho_ij = (rho_i + rho_j) / 2;
    double result = 0;
    result += -AV_ALPHA * c_s * mu_ij;
    result += AV_BETA * pow(mu_ij, 2);
    result /= rho_ij;
    return result;
}

static double pressure_of(const orc_sim *s, int j, double c_s) {
    if (s->pressure_calc == ORC_ISOTHERMAL) return pow(c_s, 2) * s->density[j];
    return (GAMMA - 1) * s->u[j] * s->density[j];
}

static double sound_speed(const orc_sim *s, int i) {
    if (s->pressure_calc == ORC_ISOTHERMAL) return 1;
    return sqrt(GAMMA * s->pressure[i] / s->density[i]);
}

static int check_density(orc_sim *s, int j) {
    if (s->density[j] < CALC_EPSILON)
        return fail(s, ORC_ERR_LOW_DENSITY, "particle had density less than epsilon", j, s->density[j]);
    return 0;
}

/* The acceleration loop visits every j != i; only candidates inside the union
 * support of h_i, h_j, (h_i+h_j)/2 can contribute, everything else adds an
 * exact zero.  The reference's density guard covers every particle, so it is
 * applied to all of them once per pass (check_all_densities). */
static int force_candidates(orc_sim *s, int i, int **out) {
    double hm = s->hmax > s->h[i] ? s->hmax : s->h[i];
    return gather(s, i, KERNEL_RADIUS * hm, out);
}

/* AccelerationCalculator::operator() — calculators.cpp:63-129.
 * omega[] holds calc_omega of every particle for its CURRENT h, density
 * (the reference re-evaluates it per pair, calculators.cpp:117; the value
 * only depends on positions and that particle's stored h, density). */
static int acceleration_one(orc_sim *s, int i, int is_ghost) {
    if (is_ghost) return 0;
    double c_s, Pr_i;
    if (s->pressure_calc == ORC_ISOTHERMAL) {
        c_s = sound_speed(s, i);
        Pr_i = pressure_of(s, i, c_s);
    } else if (s->pressure_calc == ORC_ADIABATIC) {
        Pr_i = pressure_of(s, i, 0);
        c_s = sound_speed(s, i); /* reads the STALE stored pressure */
    } else {
        return fail(s, ORC_ERR_BAD_MODE, "unknown pressure calculation mode", i, 0);
    }
    s->pressure[i] = Pr_i;
    double omega_i = s->omega[i];
    double Pr_rho_i = Pr_i / pow(s->density[i], 2) / omega_i;
    double acc = 0;
    int rc = check_density(s, i);
    if (rc) return rc;
    int *c;
    int n = force_candidates(s, i, &c);
    double xi = s->pos[i], hi = s->h[i];
    int used = 0;
    for (int k = 0; k < n; k++) {
        int j = c[k];
        if (j == i) continue;
        double r_ij = xi - s->pos[j];
        double hj = s->h[j];
        double hbig = hi > hj ? hi : hj;
        if (fabs(r_ij) > KERNEL_RADIUS * hbig * WIN_MARGIN) continue; /* all three gradients are 0 */
        used++;
        double h_ij = (hi + hj) / 2;
        double grad_W_i = grad_w(r_ij, hi);
        double grad_W_j = grad_w(r_ij, hj);
        double grad_W_ij = grad_w(r_ij, h_ij);
        double Pr_j = pressure_of(s, j, c_s);
        double omega_j = s->omega[j];
        double Pr_rho_j = Pr_j / pow(s->density[j], 2) / omega_j;
        double visc_ij = viscosity(s->vel[i], s->vel[j], s->density[i], s->density[j], r_ij, h_ij, c_s);
        double to_add = -s->mass[j] * ((grad_W_i * Pr_rho_i) + (grad_W_j * Pr_rho_j) + (grad_W_ij * visc_ij));
        acc += to_add;
    }
    s->force_nbr_sum += used;
    s->force_nbr_n++;
    s->acc[i] = acc;
    return 0;
}

/* EnergyCalculator::operator() — calculators.cpp:192-213 (no ghost check) */
static int energy_one(orc_sim *s, int i) {
    double omega = s->omega[i];
    double Pr_rho = s->pressure[i] / (omega * pow(s->density[i], 2));
    double sum = 0;
    int *c;
    int n = force_candidates(s, i, &c);
    double xi = s->pos[i], hi = s->h[i];
    for (int k = 0; k < n; k++) {
        int j = c[k];
        double r_ij = xi - s->pos[j];
        double hj = s->h[j];
        double h_ij = (hi + hj) / 2;
        if (fabs(r_ij) > KERNEL_RADIUS * h_ij * WIN_MARGIN) continue;
        double v_ij = s->vel[i] - s->vel[j];
        double c_s = sound_speed(s, i); /* fresh pressure */
        double visc = viscosity(s->vel[i], s->vel[j], s->density[i], s->density[j], r_ij, h_ij, c_s);
        double gw = grad_w(r_ij, h_ij);
        sum += Pr_rho * s->mass[j] * v_ij * gw;
        sum += 0.5 * s->mass[j] * v_ij * visc * gw;
    }
    s->du_dt[i] = sum;
    return 0;
}

static int check_all_densities(orc_sim *s) {
    for (int j = 0; j < s->n_part; j++) {
        int rc = check_density(s, j);
        if (rc) return rc;
    }
    return 0;
}

static void refresh_all_omega(orc_sim *s) {
    for (int j = 0; j < s->n_part; j++) s->omega[j] = omega_of(s, j);
}

/* ------------------------------------------------------------- public API */

orc_sim *orc_create(int n_part, double mass, int pressure_calc, double limit, double v_0,
                    double h_factor, double t_i, int variable_h, double constant_h) {
    orc_sim *s = (orc_sim *)calloc(1, sizeof *s);
    if (!s) return NULL;
    s->n_part = n_part;
    s->n_ghost = 0;
    s->cfg_mass = mass;
    s->pressure_calc = pressure_calc;
    s->limit = limit;
    s->v_0 = v_0;
    s->h_factor = h_factor;
    s->t_i = t_i;
    s->variable_h = variable_h;
    s->constant_h = constant_h;
    if (ensure_cap(s, n_part + 64)) { orc_destroy(s); return NULL; }
    return s;
}

void orc_destroy(orc_sim *s) {
    if (!s) return;
    free(s->mass); free(s->pos); free(s->vel); free(s->acc); free(s->h); free(s->du_dt);
    free(s->u); free(s->density); free(s->pressure); free(s->xs); free(s->omega);
    free(s->perm); free(s->rank); free(s->alive_perm); free(s->newton_iters); free(s->nbr_count);
    free(s->scratch);
    free(s);
}

void orc_set_order(orc_sim *s, int order) { s->order = order; }
void orc_set_index_order_sums(orc_sim *s, int on) { s->index_order_sums = on; }
int orc_n_part(const orc_sim *s) { return s->n_part; }
int orc_n_ghost(const orc_sim *s) { return s->n_ghost; }
long orc_fallback_count(const orc_sim *s) { return s->fallback_count; }
const char *orc_last_error(const orc_sim *s) { return s->errmsg; }
double orc_mean_force_neighbours(const orc_sim *s) {
    return s->force_nbr_n ? s->force_nbr_sum / (double)s->force_nbr_n : 0.0;
}

int orc_set_state(orc_sim *s, int na, int ng, const double *f) {
    if (ensure_cap(s, na + ng + 64)) return ORC_ERR_ALLOC;
    s->n_part = na + ng;
    s->n_ghost = ng;
    for (int i = 0; i < na + ng; i++) {
        const double *r = f + 9 * (size_t)i;
        s->mass[i] = r[0]; s->pos[i] = r[1]; s->vel[i] = r[2]; s->acc[i] = r[3]; s->h[i] = r[4];
        s->du_dt[i] = r[5]; s->u[i] = r[6]; s->density[i] = r[7]; s->pressure[i] = r[8];
    }
    s->alive_perm_n = -1;
    return 0;
}

void orc_get_state(const orc_sim *s, double *f, int *type) {
    int na = s->n_part - s->n_ghost;
    for (int i = 0; i < s->n_part; i++) {
        double *r = f + 9 * (size_t)i;
        r[0] = s->mass[i]; r[1] = s->pos[i]; r[2] = s->vel[i]; r[3] = s->acc[i]; r[4] = s->h[i];
        r[5] = s->du_dt[i]; r[6] = s->u[i]; r[7] = s->density[i]; r[8] = s->pressure[i];
        if (type) type[i] = i >= na;
    }
}

void orc_get_newton_iters(const orc_sim *s, int *out) {
    memcpy(out, s->newton_iters, sizeof(int) * (size_t)s->n_part);
}
void orc_get_neighbour_counts(const orc_sim *s, int *out) {
    memcpy(out, s->nbr_count, sizeof(int) * (size_t)s->n_part);
}

/* setup_ghost_particles — ghost_particles.cpp:19-115.  Ghosts are complete
 * copies of the selected alive particles (stale h, density, u, pressure,
 * acc, du_dt) with mirrored position and negated velocity, appended in
 * ascending alive index. */
int orc_ghosts(orc_sim *s) {
    int na = n_alive(s);
    double radius = KERNEL_RADIUS;
    /* first pass: count, to size the arrays */
    int g = 0;
    for (int i = 0; i < na; i++) {
        if (s->h[i] < CALC_EPSILON)
            return fail(s, ORC_ERR_GHOST_ZERO_H, "ghost particle initialization: zero smoothing length", i, s->h[i]);
        double max_pos = s->limit - (radius * s->h[i]);
        double min_pos = -s->limit + (radius * s->h[i]);
        if (s->pos[i] > max_pos || s->pos[i] < min_pos) g++;
    }
    if (ensure_cap(s, na + g + 64)) return ORC_ERR_ALLOC;
    int k = na;
    for (int i = 0; i < na; i++) {
        double max_pos = s->limit - (radius * s->h[i]);
        double min_pos = -s->limit + (radius * s->h[i]);
        if (s->pos[i] > max_pos || s->pos[i] < min_pos) {
            double vec;
            double p = s->pos[i];
            if (p < 0) vec = -s->limit - p;
            else vec = s->limit - p;
            s->mass[k] = s->mass[i];
            s->pos[k] = p + 2 * vec;
            s->vel[k] = s->vel[i] * -1;
            s->acc[k] = s->acc[i];
            s->h[k] = s->h[i];
            s->du_dt[k] = s->du_dt[i];
            s->u[k] = s->u[i];
            s->density[k] = s->density[i];
            s->pressure[k] = s->pressure[i];
            k++;
        }
    }
    s->n_ghost = g;
    s->n_part = na + g;
    return 0;
}

int orc_density(orc_sim *s, int i0, int i1) {
    rebuild_sorted(s);
    for (int i = 0; i < s->n_part; i++) { s->newton_iters[i] = 0; s->nbr_count[i] = 0; }
    for (int i = i0; i < i1; i++) {
        int rc = density_one(s, i);
        if (rc) return rc;
    }
    return 0;
}

int orc_omega(orc_sim *s, double *out) {
    rebuild_sorted(s);
    for (int i = 0; i < s->n_part; i++) out[i] = omega_of(s, i);
    return 0;
}

int orc_forces(orc_sim *s, int i0, int i1) {
    rebuild_sorted(s);
    refresh_all_omega(s);
    int rc = check_all_densities(s);
    if (rc) return rc;
    int na = n_alive(s);
    s->force_nbr_sum = 0; s->force_nbr_n = 0;
    for (int i = i0; i < i1; i++) {
        rc = acceleration_one(s, i, i >= na);
        if (rc) return rc;
        rc = energy_one(s, i);
        if (rc) return rc;
    }
    return 0;
}

/* first loop of SPHSimulation::step_forward — sph_simulation.cpp:59-73 */
int orc_kick_drift(orc_sim *s) {
    int na = n_alive(s);
    double timestep = s->t_i;
    for (int i = 0; i < na; i++) {
        s->vel[i] += s->acc[i] * (timestep / 2);
        s->u[i] += s->du_dt[i] * (timestep / 2);
        s->pos[i] += s->vel[i] * (timestep);
    }
    return 0;
}

/* SPHSimulation::step_forward — sph_simulation.cpp:55-108 */
static int step_once(orc_sim *s) {
    int rc;
    orc_kick_drift(s);
    rc = orc_ghosts(s);
    if (rc) return rc;
    rebuild_sorted(s);
    int na = n_alive(s);
    double timestep = s->t_i;
    for (int i = 0; i < s->n_part; i++) { s->newton_iters[i] = 0; s->nbr_count[i] = 0; }
    s->force_nbr_sum = 0; s->force_nbr_n = 0;

    if (s->order == ORC_ORDER_JACOBI) {
        /* all densities, then all forces from the half-kicked state, then all kicks */
        for (int i = 0; i < na; i++) { rc = density_one(s, i); if (rc) return rc; }
        refresh_all_omega(s);
        rc = check_all_densities(s);
        if (rc) return rc;
        for (int i = 0; i < na; i++) {
            rc = acceleration_one(s, i, 0); if (rc) return rc;
            rc = energy_one(s, i); if (rc) return rc;
        }
        for (int i = 0; i < na; i++) {
            s->vel[i] += s->acc[i] * (timestep / 2);
            s->u[i] += s->du_dt[i] * (timestep / 2);
        }
        return 0;
    }

    /* reference order (sph_simulation.cpp:94-107): in array order and IN PLACE, so
     * particle i sees new h, rho, fully-kicked v, u for every j < i and last step's
     * for every j > i.  omega[] starts as "stale" for everyone
     * (old h, rho at the new positions and ghost set) and entry i is refreshed as
     * soon as dc(p_i) has run. */
    refresh_all_omega(s);
    rc = check_all_densities(s);
    if (rc) return rc;
    for (int i = 0; i < na; i++) {
        rc = density_one(s, i); if (rc) return rc;
        rc = check_density(s, i); if (rc) return rc;
        s->omega[i] = omega_of(s, i);
        rc = acceleration_one(s, i, 0); if (rc) return rc;
        rc = energy_one(s, i); if (rc) return rc;
        s->vel[i] += s->acc[i] * (timestep / 2);
        s->u[i] += s->du_dt[i] * (timestep / 2);
    }
    return 0;
}

int orc_step(orc_sim *s, int nsteps) {
    for (int k = 0; k < nsteps; k++) {
        int rc = step_once(s);
        if (rc) return rc;
    }
    return 0;
}

/* init_particles — setup.cpp:143-228 */
int orc_init_particles(orc_sim *s) {
    int n = s->n_part - s->n_ghost;
    s->n_part = n;
    s->n_ghost = 0;
    double max_x = s->limit;
    double min_x = -max_x;
    double spacing = (max_x - min_x) / (n - 1);
    max_x -= spacing / 2;
    min_x += spacing / 2;
    for (int i = 0; i < n; i++) {
        double spacing2 = (max_x - min_x) / (n - 1);
        double pos = min_x + spacing2 * (i);
        double vel = (pos < 0) ? s->v_0 : -s->v_0;
        s->pos[i] = pos;
        s->mass[i] = s->cfg_mass;
        s->vel[i] = vel;
        s->acc[i] = 0; s->du_dt[i] = 0; s->density[i] = 0; s->pressure[i] = 0;
        s->u[i] = 0; /* the reference leaves u uninitialised in isothermal runs (setup.cpp:171-172) */
        if (s->pressure_calc == ORC_ADIABATIC) s->u[i] = 1 / (GAMMA - 1);
        if (s->variable_h) s->h[i] = s->h_factor * spacing2;
        else s->h[i] = s->constant_h;
    }
    s->alive_perm_n = -1;
    int rc;
    if (s->pressure_calc == ORC_ADIABATIC) {
        /* pre-pass, setup.cpp:192-202: densities without ghosts; then, in array order
         * and in place, ac(p_i) followed by v_i = +-c_s from the pressure ac() just
         * stored.  That stored pressure is the stale pressure the t = 0 force pass
         * reads; the pre-pass acceleration itself only survives in
         * the ghost copies made below (alive values are overwritten at t = 0). */
        rc = orc_density(s, 0, n);
        if (rc) return rc;
        rc = check_all_densities(s);
        if (rc) return rc;
        refresh_all_omega(s);
        for (int i = 0; i < n; i++) {
            rc = acceleration_one(s, i, 0);
            if (rc) return rc;
            double c_s = sound_speed(s, i);
            s->vel[i] = (s->pos[i] < 0) ? c_s : -c_s;
        }
    }
    rc = orc_ghosts(s);
    if (rc) return rc;
    rc = orc_density(s, 0, s->n_part);
    if (rc) return rc;
    return orc_forces(s, 0, s->n_part);
}
