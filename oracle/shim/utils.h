// oracle/shim/utils.h -- TEST INFRASTRUCTURE ONLY.  Restatement of the reference's missing `utils.h`.
// Contracts inferred from call sites: alloc_2d_array<T>(n,m) returns T** whose rows live in ONE contiguous
// n*m block starting at a[0] (iteration_lvg.cpp:55,117 memsets *matrix as a block); free_2d_array releases it;
// rounding(x) = nearest integer (spectroscopy.h:58); locate_index is in interpolation.h but dust/photoelectric
// code reaches it through utils.h, so it is declared here and included from there.
#pragma once
#include <cmath>
#include <cstddef>
#include <cfloat>
#include <iomanip>   // the original utils.h evidently pulls these in (dust_model.cpp:881 uses setw, photoelectric_emission.cpp:389 DBL_EPSILON)
#include <iostream>

template <class T> T** alloc_2d_array(int n, int m)
{
    T** a = new T*[n > 0 ? n : 1];
    a[0] = new T[(size_t)(n > 0 ? n : 1) * (size_t)(m > 0 ? m : 1)];
    for (int i = 1; i < n; i++) a[i] = a[0] + (size_t)i * m;
    return a;
}

template <class T> void free_2d_array(T** a)
{
    if (a) { delete[] a[0]; delete[] a; }
}

inline int rounding(double x) { return (int)std::floor(x + 0.5); }

// Ascending grid arr[0..n-1]; returns k with arr[k] <= x < arr[k+1]; k = -1 below the grid, k = n-1 at or
// above the last node (callers clamp: lvg_method_functions.cpp:82-100).
inline void locate_index(const double* arr, int n, double x, int& k)
{
    if (!(x >= arr[0])) { k = -1; return; }
    if (x >= arr[n - 1]) { k = n - 1; return; }
    int lo = 0, hi = n - 1;
    while (hi - lo > 1) {
        int mid = lo + ((hi - lo) >> 1);
        if (arr[mid] <= x) lo = mid; else hi = mid;
    }
    k = lo;
}
