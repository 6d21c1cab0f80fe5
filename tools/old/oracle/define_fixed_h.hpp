/*
 * oracle/define_fixed_h.hpp — TEST INFRASTRUCTURE, not product code.
 *
 * The reference selects fixed vs variable smoothing length at COMPILE time
 * (reference sph/define.hpp:27 "#define USE_VARIABLE_H", :29 CONSTANT_H = 0.2).
 * Its sources must not be edited, so the fixed-h oracle build force-includes
 * this file (g++ -include): it pre-defines the reference header's include
 * guard, which empties reference sph/define.hpp, and restates that header's
 * settings with USE_VARIABLE_H left undefined and CONSTANT_H redirected to a
 * run-time variable owned by oracle/ref_harness.cpp (so BASELINE config 2's
 * resolution-scaled h can be used).
 * Values restated from reference sph/define.hpp:25,32,39,41,43,46.
 */
#ifndef SPHB_ORACLE_DEFINE_FIXED_H
#define SPHB_ORACLE_DEFINE_FIXED_H
#define define_hpp
#define SPHB_ORACLE_FIXED_H 1

const double CALC_EPSILON = 1e-8;
extern "C" double sphb_oracle_constant_h;
#define CONSTANT_H sphb_oracle_constant_h
const double GAMMA = 5.0 / 3.0;
#define H_MAX_ITER_NR 10
#define H_MAX_ITER_BS 1000
#define H_EPSILON 1e-4
#define H_WARNINGS

#endif
