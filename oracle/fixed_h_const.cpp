/*
 * oracle/fixed_h_const.cpp — TEST INFRASTRUCTURE, not product code.
 * The value behind CONSTANT_H for the stand-alone fixed-h build of the reference program
 * (oracle/_ref/sph_ref_fixedh): reference sph/define.hpp:29.  See define_fixed_h.hpp.
 */
extern "C" double sphb_oracle_constant_h;
double sphb_oracle_constant_h = 0.2;
