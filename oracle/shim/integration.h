// oracle/shim/integration.h -- TEST INFRASTRUCTURE ONLY.  Restatement of the reference's missing header.
// qromb<F>(f, a, b, eps): Romberg integration of f over [a,b] to relative accuracy eps (call sites:
// lvg_method_functions.cpp:129, dust_model.cpp:250-253,551).  Set-up-time only; never on the hot path.
#pragma once
#include <cmath>

template <class F> double qromb(F& f, double a, double b, double eps)
{
    const int JMAX = 22;
    double R[JMAX + 1][JMAX + 1];
    double h = b - a;
    R[0][0] = 0.5 * h * (f(a) + f(b));
    for (int j = 1; j <= JMAX; j++) {
        h *= 0.5;
        double s = 0.;
        const long n = 1L << (j - 1);
        for (long i = 0; i < n; i++) s += f(a + (2 * i + 1) * h);
        R[j][0] = 0.5 * R[j - 1][0] + h * s;
        double p4 = 1.;
        for (int m = 1; m <= j; m++) {
            p4 *= 4.;
            R[j][m] = R[j][m - 1] + (R[j][m - 1] - R[j - 1][m - 1]) / (p4 - 1.);
        }
        if (j >= 4 && std::fabs(R[j][j] - R[j - 1][j - 1]) <= eps * std::fabs(R[j][j])) return R[j][j];
    }
    return R[JMAX][JMAX];
}
