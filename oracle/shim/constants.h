// oracle/shim/constants.h -- TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Restatement of the reference's missing `constants.h` (pulled from the author's private
// `input_data/mathematics/` directory, radiative_transfer.vcxproj:93).  CGS units; CODATA-2014-ish values.
// Names are fixed by their call sites in the reference (e.g. iteration_lvg.cpp:65,168,455; coll_rates.cpp:194;
// transition_data.cpp:234,268).  Values are the builder's: "parity unpinned" at this boundary -- the same
// numbers are used by the CUDA path (include/rt_b200.h: rt_params) so both sides agree by construction.
#pragma once
#ifndef _USE_MATH_DEFINES
#define _USE_MATH_DEFINES
#endif
#include <cmath>

#define SPEED_OF_LIGHT        2.99792458e+10      // cm/s
#define BOLTZMANN_CONSTANT    1.38064852e-16      // erg/K
#define PLANCK_CONSTANT       6.62607004e-27      // erg s
#define ATOMIC_MASS_UNIT      1.660539040e-24     // g
#define ELECTRON_CHARGE       4.80320451e-10      // statcoulomb
#define DEBYE                 1.e-18              // statC cm
#define EV_TO_ERGS            1.6021766208e-12
#define CM_INVERSE_TO_ERG     (PLANCK_CONSTANT*SPEED_OF_LIGHT)
#define ERG_TO_CM_INVERSE     (1./(PLANCK_CONSTANT*SPEED_OF_LIGHT))
#define CM_INVERSE_TO_KELVINS 1.438777364         // hc/k, K per cm-1
#define CM_INVERSE_TO_EV      (PLANCK_CONSTANT*SPEED_OF_LIGHT/EV_TO_ERGS)
#define EV_TO_CM_INVERSE      (EV_TO_ERGS/(PLANCK_CONSTANT*SPEED_OF_LIGHT))
#define EIGHT_PI              25.132741228718345
#define SQRT_PI               1.7724538509055160
#define ONEDIVBY_SQRT_PI      0.56418958354775629
#define EXT_MAG_TO_OPT_DEPTH  0.92103403719761834  // ln(10)/2.5 = 1/1.0857
