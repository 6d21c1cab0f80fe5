/*
 * oracle/gsl_shim/gsl/gsl_roots.h — TEST INFRASTRUCTURE, not product code.
 *
 * The reference links the system GNU Scientific Library (un-pinned, un-vendored:
 * reference sph/Makefile:5 "-lgsl -lgslcblas") for exactly two root solvers,
 * called from reference sph/smoothing_length.cpp:165-185 (bisection) and
 * :232-257 (Newton).  GSL is not installed in this image, so this header
 * restates the published behaviour of those two solvers and of the two
 * convergence tests (GSL manual, "One Dimensional Root-Finding"; SURVEY.md
 * Appendix B) with the same names and signatures, so that the reference's
 * sources compile UNMODIFIED against it.  Nothing here is taken from the
 * reference tree.  Parity at this boundary is unpinned by the reference's own
 * tests; see DESIGN.md.
 *
 * Extra, for the parity tests only: every Newton root-find appends its
 * iteration count (and whether the bisection fallback then ran) to a log
 * that the harness reads back (oracle/ref_harness.cpp).
 */
#ifndef SPHB_ORACLE_GSL_ROOTS_SHIM_H
#define SPHB_ORACLE_GSL_ROOTS_SHIM_H

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "gsl_errno.h"

/* ---- callback records, as the GSL manual documents them ---- */
struct gsl_function {
    double (*function)(double x, void *params);
    void *params;
};

struct gsl_function_fdf {
    double (*f)(double x, void *params);
    double (*df)(double x, void *params);
    void (*fdf)(double x, void *params, double *f, double *df);
    void *params;
};

/* ---- instrumentation shared by all translation units (C++17 inline vars) ---- */
struct sphb_shim_log_t {
    std::vector<int> newton_iters;   /* one entry per fdfsolver lifetime (alloc..free) */
    std::vector<int> bisect_iters;   /* one entry per fsolver lifetime */
    long fdf_evals = 0;
};
inline sphb_shim_log_t sphb_shim_log;

[[noreturn]] inline void sphb_shim_abort(const char *reason, int code) {
    /* GSL's default error handler prints and aborts the process. */
    std::fprintf(stderr, "gsl (shim): ERROR: %s (errno %d)\nDefault GSL error handler invoked.\n",
                 reason, code);
    std::abort();
}

/* =========================== bracketing solver ============================ */
struct gsl_root_fsolver_type { const char *name; };
inline const gsl_root_fsolver_type sphb_shim_bisection_type = {"bisection"};
inline const gsl_root_fsolver_type *gsl_root_fsolver_bisection = &sphb_shim_bisection_type;

struct gsl_root_fsolver {
    const gsl_root_fsolver_type *type;
    gsl_function *function;
    double root, x_lower, x_upper;
    double f_lower, f_upper;
    int iters;
};

inline gsl_root_fsolver *gsl_root_fsolver_alloc(const gsl_root_fsolver_type *T) {
    gsl_root_fsolver *s = new gsl_root_fsolver();
    s->type = T;
    s->function = nullptr;
    s->iters = 0;
    return s;
}

inline double sphb_shim_safe_eval(gsl_function *f, double x) {
    double y = f->function(x, f->params);
    if (!std::isfinite(y)) sphb_shim_abort("function value is not finite", GSL_EBADFUNC);
    return y;
}

inline int gsl_root_fsolver_set(gsl_root_fsolver *s, gsl_function *f, double lo, double hi) {
    if (lo > hi) sphb_shim_abort("invalid interval (lower > upper)", GSL_EINVAL);
    s->function = f;
    s->x_lower = lo;
    s->x_upper = hi;
    s->root = 0.5 * (lo + hi);
    s->f_lower = sphb_shim_safe_eval(f, lo);
    s->f_upper = sphb_shim_safe_eval(f, hi);
    if ((s->f_lower < 0.0 && s->f_upper < 0.0) || (s->f_lower > 0.0 && s->f_upper > 0.0))
        sphb_shim_abort("endpoints do not straddle y=0", GSL_EINVAL);
    return GSL_SUCCESS;
}

inline int gsl_root_fsolver_iterate(gsl_root_fsolver *s) {
    s->iters++;
    const double left = s->x_lower, right = s->x_upper;
    if (s->f_lower == 0.0) { s->root = left;  s->x_upper = left;  return GSL_SUCCESS; }
    if (s->f_upper == 0.0) { s->root = right; s->x_lower = right; return GSL_SUCCESS; }
    const double mid = (left + right) / 2.0;
    const double f_mid = sphb_shim_safe_eval(s->function, mid);
    if (f_mid == 0.0) {
        s->root = mid; s->x_lower = mid; s->x_upper = mid;
        return GSL_SUCCESS;
    }
    const bool sign_change_left =
        (s->f_lower > 0.0 && f_mid < 0.0) || (s->f_lower < 0.0 && f_mid > 0.0);
    if (sign_change_left) {
        s->root = 0.5 * (left + mid);
        s->x_upper = mid;
        s->f_upper = f_mid;
    } else {
        s->root = 0.5 * (mid + right);
        s->x_lower = mid;
        s->f_lower = f_mid;
    }
    return GSL_SUCCESS;
}

inline double gsl_root_fsolver_root(const gsl_root_fsolver *s) { return s->root; }
inline double gsl_root_fsolver_x_lower(const gsl_root_fsolver *s) { return s->x_lower; }
inline double gsl_root_fsolver_x_upper(const gsl_root_fsolver *s) { return s->x_upper; }

inline void gsl_root_fsolver_free(gsl_root_fsolver *s) {
    if (!s) return;
    sphb_shim_log.bisect_iters.push_back(s->iters);
    delete s;
}

/* ============================ derivative solver =========================== */
struct gsl_root_fdfsolver_type { const char *name; };
inline const gsl_root_fdfsolver_type sphb_shim_newton_type = {"newton"};
inline const gsl_root_fdfsolver_type *gsl_root_fdfsolver_newton = &sphb_shim_newton_type;

struct gsl_root_fdfsolver {
    const gsl_root_fdfsolver_type *type;
    gsl_function_fdf *fdf;
    double root;
    double f, df;
    int iters;
};

inline gsl_root_fdfsolver *gsl_root_fdfsolver_alloc(const gsl_root_fdfsolver_type *T) {
    gsl_root_fdfsolver *s = new gsl_root_fdfsolver();
    s->type = T;
    s->fdf = nullptr;
    s->iters = 0;
    return s;
}

inline int gsl_root_fdfsolver_set(gsl_root_fdfsolver *s, gsl_function_fdf *fdf, double root) {
    s->fdf = fdf;
    s->root = root;
    fdf->fdf(root, fdf->params, &s->f, &s->df);
    sphb_shim_log.fdf_evals++;
    return GSL_SUCCESS;
}

inline int gsl_root_fdfsolver_iterate(gsl_root_fdfsolver *s) {
    s->iters++;
    if (s->df == 0.0) sphb_shim_abort("derivative is zero", GSL_EZERODIV);
    const double next = s->root - (s->f / s->df);
    s->root = next;
    double f_new, df_new;
    s->fdf->fdf(next, s->fdf->params, &f_new, &df_new);
    sphb_shim_log.fdf_evals++;
    s->f = f_new;
    s->df = df_new;
    if (!std::isfinite(f_new)) sphb_shim_abort("function value is not finite", GSL_EBADFUNC);
    if (!std::isfinite(df_new)) sphb_shim_abort("derivative value is not finite", GSL_EBADFUNC);
    return GSL_SUCCESS;
}

inline double gsl_root_fdfsolver_root(const gsl_root_fdfsolver *s) { return s->root; }

inline void gsl_root_fdfsolver_free(gsl_root_fdfsolver *s) {
    if (!s) return;
    sphb_shim_log.newton_iters.push_back(s->iters);
    delete s;
}

/* ============================ convergence tests =========================== */
inline int gsl_root_test_delta(double x1, double x0, double epsabs, double epsrel) {
    if (epsabs < 0.0) sphb_shim_abort("absolute tolerance is negative", GSL_EBADTOL);
    if (epsrel < 0.0) sphb_shim_abort("relative tolerance is negative", GSL_EBADTOL);
    const double tol = epsabs + epsrel * std::fabs(x1);
    if (std::fabs(x1 - x0) < tol || x1 == x0) return GSL_SUCCESS;
    return GSL_CONTINUE;
}

inline int gsl_root_test_interval(double lo, double hi, double epsabs, double epsrel) {
    if (epsabs < 0.0) sphb_shim_abort("absolute tolerance is negative", GSL_EBADTOL);
    if (epsrel < 0.0) sphb_shim_abort("relative tolerance is negative", GSL_EBADTOL);
    if (lo > hi) sphb_shim_abort("lower bound larger than upper bound", GSL_EINVAL);
    const double alo = std::fabs(lo), ahi = std::fabs(hi);
    double smallest = 0.0;
    if ((lo > 0.0 && hi > 0.0) || (lo < 0.0 && hi < 0.0)) smallest = (alo < ahi) ? alo : ahi;
    const double tol = epsabs + epsrel * smallest;
    if (std::fabs(hi - lo) < tol) return GSL_SUCCESS;
    return GSL_CONTINUE;
}

#endif
