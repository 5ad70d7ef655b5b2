// oracle/shim/interpolation.h -- TEST INFRASTRUCTURE ONLY.  Restatement of the reference's missing header.
// locate_index lives in utils.h here.  bcuint is only reached from lvg_method_data_spline::get_esc_func
// (lvg_method_functions.cpp:254), which has no live caller; it is provided as a bilinear stand-in so the file links.
#pragma once
#include "utils.h"

inline void bcuint(const double* y, const double*, const double*, const double*, double x1l, double x1u,
                   double x2l, double x2u, double x1, double x2, double& ansy, double& ansy1, double& ansy2)
{
    // corner numbering counter-clockwise from lower-left, as at the call site (lvg_method_functions.cpp:236-252)
    const double t = (x1 - x1l) / (x1u - x1l), u = (x2 - x2l) / (x2u - x2l);
    ansy = (1. - t) * (1. - u) * y[0] + t * (1. - u) * y[1] + t * u * y[2] + (1. - t) * u * y[3];
    ansy1 = ansy2 = 0.;
}
