// oracle/shim/linear_algebra.h -- TEST INFRASTRUCTURE ONLY.
// Restatement of the reference's missing `linear_algebra.h`.  The only symbol the hot path uses is
//     lu_matrix_solve(double** a, double* b, int n)
// (call sites: iteration_lvg.cpp:100, iteration_control.h:176, iteration_control.cpp:87, coll_rates.cpp:396).
// Contract inferred from those sites: `a` is a row-pointer matrix (rows contiguous), it is destroyed; `b` holds
// the right-hand side on entry and the solution on return.  The original's pivoting strategy is UNKNOWN
// ("parity unpinned"); this restatement is the textbook dense LU with row partial pivoting (largest |a_ik| in
// the column, first occurrence wins ties), right-looking row-oriented elimination, followed by forward and
// back substitution -- the dgesv contract.  The CUDA solver (csrc/lu_*.cu) implements the same pivot rule.
#pragma once
#include <cmath>
#include <utility>

inline void lu_matrix_solve(double** a, double* b, int n)
{
    for (int k = 0; k < n; k++) {
        int p = k;
        double big = std::fabs(a[k][k]);
        for (int i = k + 1; i < n; i++) {
            double v = std::fabs(a[i][k]);
            if (v > big) { big = v; p = i; }
        }
        if (p != k) {
            std::swap(a[p], a[k]);          // row pointers: O(1) interchange
            std::swap(b[p], b[k]);
        }
        const double piv = a[k][k];
        if (piv == 0.) continue;            // singular column: leave as is (solution will hold inf/nan)
        const double rp = 1. / piv;
        const double* rk = a[k];
        for (int i = k + 1; i < n; i++) {
            double* ri = a[i];
            const double l = ri[k] * rp;
            if (l != 0.) {
                ri[k] = l;
                for (int j = k + 1; j < n; j++) ri[j] -= l * rk[j];
                b[i] -= l * b[k];
            }
        }
    }
    for (int i = n - 1; i >= 0; i--) {
        double s = b[i];
        const double* ri = a[i];
        for (int j = i + 1; j < n; j++) s -= ri[j] * b[j];
        b[i] = s / ri[i];
    }
    // NB: row pointers were permuted; alloc_2d_array's a[0] may no longer be the block start.  Restore so that
    // free_2d_array(a) and `memset(*matrix, ...)` (iteration_lvg.cpp:117) keep working: sort pointers ascending.
    for (int i = 1; i < n; i++) {
        double* r = a[i]; int j = i - 1;
        while (j >= 0 && a[j] > r) { a[j + 1] = a[j]; j--; }
        a[j + 1] = r;
    }
}
