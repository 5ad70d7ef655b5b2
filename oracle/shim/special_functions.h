// oracle/shim/special_functions.h -- TEST INFRASTRUCTURE ONLY.  Restatement of the reference's missing header.
// expint(n, x): exponential integral E_n(x) (lvg_method_functions.cpp:118,125; dead path get_esc_func_g).
// err_func: functor with member f(x) = erf(x) (dust_model.cpp:87-96, set-up only).
#pragma once
#include <cmath>
#include "utils.h"

struct err_func {
    double f(double x) const { return std::erf(x); }
    double operator()(double x) const { return std::erf(x); }
};

inline double expint(int n, double x)
{
    // continued fraction for x > 1, series otherwise (Abramowitz & Stegun 5.1.11-5.1.12, 5.1.22)
    const double EULER = 0.57721566490153286, EPS = 1e-15, FPMIN = 1e-300;
    const int nm1 = n - 1;
    if (x == 0.) return nm1 > 0 ? 1. / nm1 : HUGE_VAL;
    if (x > 1.) {
        double b = x + n, c = 1. / FPMIN, d = 1. / b, h = d;
        for (int i = 1; i <= 10000; i++) {
            const double an = -i * (double)(nm1 + i);
            b += 2.; d = 1. / (an * d + b); c = b + an / c;
            const double del = c * d; h *= del;
            if (std::fabs(del - 1.) < EPS) break;
        }
        return h * std::exp(-x);
    }
    double ans = nm1 != 0 ? 1. / nm1 : -std::log(x) - EULER, fact = 1.;
    for (int i = 1; i <= 10000; i++) {
        fact *= -x / i;
        double del;
        if (i != nm1) del = -fact / (i - nm1);
        else {
            double psi = -EULER;
            for (int ii = 1; ii <= nm1; ii++) psi += 1. / ii;
            del = fact * (-std::log(x) + psi);
        }
        ans += del;
        if (std::fabs(del) < std::fabs(ans) * EPS) break;
    }
    return ans;
}
