/*
 * oracle/gsl_shim/gsl/gsl_errno.h — TEST INFRASTRUCTURE (see gsl_roots.h).
 * Status codes and gsl_strerror as documented in the GSL manual ("Error
 * Codes"); only the ones the two root solvers can produce are named.
 */
#ifndef SPHB_ORACLE_GSL_ERRNO_SHIM_H
#define SPHB_ORACLE_GSL_ERRNO_SHIM_H

enum {
    GSL_SUCCESS = 0,
    GSL_FAILURE = -1,
    GSL_CONTINUE = -2,
    GSL_EDOM = 1,
    GSL_ERANGE = 2,
    GSL_EFAULT = 3,
    GSL_EINVAL = 4,
    GSL_EFAILED = 5,
    GSL_EBADFUNC = 9,
    GSL_EMAXITER = 11,
    GSL_EZERODIV = 12,
    GSL_EBADTOL = 13
};

inline const char *gsl_strerror(const int code) {
    switch (code) {
        case GSL_SUCCESS: return "success";
        case GSL_FAILURE: return "failure";
        case GSL_CONTINUE: return "the iteration has not converged yet";
        case GSL_EDOM: return "input domain error";
        case GSL_ERANGE: return "output range error";
        case GSL_EFAULT: return "invalid pointer";
        case GSL_EINVAL: return "invalid argument supplied by user";
        case GSL_EFAILED: return "generic failure";
        case GSL_EBADFUNC: return "problem with user-supplied function";
        case GSL_EMAXITER: return "exceeded max number of iterations";
        case GSL_EZERODIV: return "tried to divide by zero";
        case GSL_EBADTOL: return "specified tolerance is invalid or theoretically unattainable";
        default: return "unknown error code";
    }
}
#endif
