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

// Rounding-sensitivity probe (test infrastructure, used ONLY when tests/golden/make_full_slab.py maps which layers of a
// slab are sensitive to the rounding of this solve; the default 0 is the textbook solve every fixture answer comes from):
//   bit 0: every multiply-subtract of the elimination and of the back substitution is ONE fused operation (one rounding
//          instead of two) -- same algorithm, same order of operations;
//   bit 1: every entry of the matrix is moved by one unit in the last place up or down (a fixed pattern) before the
//          elimination, as a different summation order of the diagonal sums in the assembly would;
//   bit 2: the other half of that pattern.
// A layer whose answer moves by more than the parity tolerance under these probes has an answer that is decided by
// rounding, in the reference itself.  Only systems larger than 8 x 8 are touched (not the 4 x 4 solve of the Ng step).
inline int& shim_lu_variant() { static int v = 0; return v; }

template <bool FUSED>
__attribute__((always_inline)) inline void lu_matrix_solve_body(double** a, double* b, int n)
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
                if (FUSED) {
                    for (int j = k + 1; j < n; j++) ri[j] = __builtin_fma(-l, rk[j], ri[j]);
                    b[i] = __builtin_fma(-l, b[k], b[i]);
                }
                else {
                    for (int j = k + 1; j < n; j++) ri[j] -= l * rk[j];
                    b[i] -= l * b[k];
                }
            }
        }
    }
    for (int i = n - 1; i >= 0; i--) {
        double s = b[i];
        const double* ri = a[i];
        if (FUSED) for (int j = i + 1; j < n; j++) s = __builtin_fma(-ri[j], b[j], s);
        else for (int j = i + 1; j < n; j++) s -= ri[j] * b[j];
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

__attribute__((target("fma"))) inline void lu_matrix_solve_fused(double** a, double* b, int n) { lu_matrix_solve_body<true>(a, b, n); }

inline void lu_matrix_solve(double** a, double* b, int n)
{
    const int variant = n > 8 ? shim_lu_variant() : 0;
    if (variant & 6) {
        for (int i = 0; i < n; i++)
            for (int j = 0; j < n; j++) {
                const unsigned h = (unsigned)(i * 2654435761u) ^ (unsigned)(j * 40503u) ^ (unsigned)((i + 3) * (j + 7));
                const unsigned sel = (h >> 7) & 3u; // a quarter of the entries up, a quarter down, per probe bit
                double& x = a[i][j];
                if ((variant & 2) && sel == 0) x = std::nextafter(x, x > 0 ? HUGE_VAL : -HUGE_VAL);
                if ((variant & 2) && sel == 1) x = std::nextafter(x, 0.);
                if ((variant & 4) && sel == 2) x = std::nextafter(x, x > 0 ? HUGE_VAL : -HUGE_VAL);
                if ((variant & 4) && sel == 3) x = std::nextafter(x, 0.);
            }
    }
    if (variant & 1) lu_matrix_solve_fused(a, b, n);
    else lu_matrix_solve_body<false>(a, b, n);
}
