// fixed_h_build.hpp — force-included (g++ -include) to build the drop-in the way the reference
// builds its fixed-smoothing-length variant: by leaving USE_VARIABLE_H undefined
// (reference sph/define.hpp:27-29).  The reference header cannot be edited, so this file
// pre-defines its include guard and restates its settings (define.hpp:25,29,32,39,41,43).
#ifndef SPHB_FIXED_H_BUILD
#define SPHB_FIXED_H_BUILD
#define define_hpp
const double CALC_EPSILON = 1e-8;
const double CONSTANT_H = 0.2;
const double GAMMA = 5.0 / 3.0;
#define H_MAX_ITER_NR 10
#define H_MAX_ITER_BS 1000
#define H_EPSILON 1e-4
#define H_WARNINGS
#endif
