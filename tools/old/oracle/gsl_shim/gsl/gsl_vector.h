/* oracle/gsl_shim/gsl/gsl_vector.h — TEST INFRASTRUCTURE (see gsl_roots.h).
 * The reference includes this header (sph/smoothing_length.cpp:1) but uses
 * nothing from it; an empty stand-in is enough. */
#ifndef SPHB_ORACLE_GSL_VECTOR_SHIM_H
#define SPHB_ORACLE_GSL_VECTOR_SHIM_H
#endif
