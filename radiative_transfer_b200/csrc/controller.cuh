// controller.cuh -- device-side restatement of iteration_control<T> (iteration_control.h:69-242) shared by the
// multi-kernel path (controller_kernels.cu) and the fused small-N kernel (fused_kernels.cu).
#pragma once
#include "rt_internal.h"

#define CTL_THREADS 256

static __device__ __forceinline__ double* ring_slot(double* base, int lay, int N, int slot)
{
    return base + ((size_t)lay * RT_HIST + slot) * N;
}

// max-reduce two non-negative values over the CTA; result valid in all threads
static __device__ __forceinline__ void block_max2(double& a, double& b)
{
    __shared__ double sa[CTL_THREADS / 32], sb[CTL_THREADS / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a = fmax(a, __shfl_xor_sync(0xffffffffu, a, o));
        b = fmax(b, __shfl_xor_sync(0xffffffffu, b, o));
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { sa[warp] = a; sb[warp] = b; }
    __syncthreads();
    a = sa[0]; b = sb[0];
#pragma unroll
    for (int w = 1; w < CTL_THREADS / 32; w++) { a = fmax(a, sa[w]); b = fmax(b, sb[w]); }
    __syncthreads();
}

// 4x4 dense solve with the pivot rule of lu_matrix_solve (oracle/shim/linear_algebra.h), single thread,
// no FMA contraction so that identical inputs give the bits of the scalar CPU code.
static __device__ void solve4(double* A, double* b)
{
    const int n = 4;
    for (int k = 0; k < n; k++) {
        int p = k;
        double big = fabs(A[k * n + k]);
        for (int i = k + 1; i < n; i++) {
            const double v = fabs(A[i * n + k]);
            if (v > big) { big = v; p = i; }
        }
        if (p != k) {
            for (int j = 0; j < n; j++) { const double t = A[p * n + j]; A[p * n + j] = A[k * n + j]; A[k * n + j] = t; }
            const double t = b[p]; b[p] = b[k]; b[k] = t;
        }
        const double piv = A[k * n + k];
        if (piv == 0.) continue;
        const double rp = __ddiv_rn(1., piv);
        for (int i = k + 1; i < n; i++) {
            const double l = __dmul_rn(A[i * n + k], rp);
            if (l != 0.) {
                for (int j = k + 1; j < n; j++) A[i * n + j] = __dsub_rn(A[i * n + j], __dmul_rn(l, A[k * n + j]));
                b[i] = __dsub_rn(b[i], __dmul_rn(l, b[k]));
            }
        }
    }
    for (int i = n - 1; i >= 0; i--) {
        double s = b[i];
        for (int j = i + 1; j < n; j++) s = __dsub_rn(s, __dmul_rn(A[i * n + j], b[j]));
        b[i] = __ddiv_rn(s, A[i * n + i]);
    }
}

// accel_step, iteration_control.h:139-193 (nb_prev_steps = 5 -> 4 unknowns).  The 16 + 4 normal-equation sums run
// over k = 0..N-1 in the reference's order, one accumulator per thread, without FMA contraction.
static __device__ void accel_step(const DevModel& d, int lay, double* pop)
{
    const int N = d.N;
    __shared__ double s_A[16], s_b[4], s_c[4];
    const int hp = d.head_pop[lay], hr = d.head_res[lay];
    const int t = threadIdx.x;
    if (t < 20) {
        const double* r0 = ring_slot(d.hist_res, lay, N, hr);
        const double* n0 = ring_slot(d.hist_pop, lay, N, hp);
        const int i = t < 16 ? t >> 2 : t - 16, j = t & 3;
        const double* ri = ring_slot(d.hist_res, lay, N, (hr + i + 1) % RT_HIST);
        const double* rj = ring_slot(d.hist_res, lay, N, (hr + j + 1) % RT_HIST);
        double acc = 0.;
        if (t < 16) {
            for (int k = 0; k < N; k++) {
                const double w = __dadd_rn(n0[k], 1.e-99);
                const double num = __dmul_rn(__dsub_rn(r0[k], ri[k]), __dsub_rn(r0[k], rj[k]));
                acc = __dadd_rn(acc, __ddiv_rn(num, __dmul_rn(w, w)));
            }
            s_A[t] = acc;
        }
        else {
            for (int k = 0; k < N; k++) {
                const double w = __dadd_rn(n0[k], 1.e-99);
                const double num = __dmul_rn(__dsub_rn(r0[k], ri[k]), r0[k]);
                acc = __dadd_rn(acc, __ddiv_rn(num, __dmul_rn(w, w)));
            }
            s_b[t - 16] = acc;
        }
    }
    __syncthreads();
    if (t == 0) {
        double A[16], b[4];
        for (int q = 0; q < 16; q++) A[q] = s_A[q];
        for (int q = 0; q < 4; q++) b[q] = s_b[q];
        solve4(A, b);
        for (int q = 0; q < 4; q++) s_c[q] = b[q];
    }
    __syncthreads();
    double sum = 0.;
    for (int q = 0; q < 4; q++) sum = __dadd_rn(sum, s_c[q]);
    const double* P[5];
    for (int q = 0; q < 5; q++) P[q] = ring_slot(d.hist_pop, lay, N, (hp + q) % RT_HIST);
    for (int k = t; k < N; k += CTL_THREADS) {
        double v = __dmul_rn(__dsub_rn(1., sum), P[0][k]);
        for (int q = 0; q < 4; q++) v = __dadd_rn(v, __dmul_rn(s_c[q], P[q + 1][k]));
        pop[k] = v;
    }
    __syncthreads();
}

// head of next_step for the iteration about to run (iteration_control.h:92-103)
static __device__ void pre_step(const DevModel& d, int lay, int accel)
{
    const int N = d.N;
    double* pop = d.pop + (size_t)lay * N;
    // prev_level_pop.push_front(pop_old)
    int hp = (d.head_pop[lay] + RT_HIST - 1) % RT_HIST;
    double* slot = ring_slot(d.hist_pop, lay, N, hp);
    for (int k = threadIdx.x; k < N; k += CTL_THREADS) slot[k] = pop[k];
    __syncthreads();
    const int iter = d.iter_nb[lay];
    int nb_after = d.nb_after[lay];
    const int n_res = d.n_hist_res[lay];
    __syncthreads();
    if (threadIdx.x == 0) {
        d.head_pop[lay] = hp;
        if (d.n_hist_pop[lay] < RT_HIST) d.n_hist_pop[lay]++;
    }
    __syncthreads();
    if (accel && (iter == d.prm.accel_start || nb_after == d.prm.accel_period) && n_res >= 5) {
        accel_step(d, lay, pop);
        nb_after = 0;
    }
    if (accel && iter >= d.prm.accel_start) nb_after++;
    if (threadIdx.x == 0) d.nb_after[lay] = nb_after;
}

// tail of next_step for iteration n (iteration_control.h:105-136) and, if the layer goes on, the head of next_step for
// iteration n+1 (:92-98).  Whole-CTA cooperative (CTL_THREADS threads).  Returns true when the layer stops.
static __device__ bool ctl_step_body(const DevModel& d, int lay, int accel, int max_iter, double min_err)
{
    const int N = d.N;
    double* pop = d.pop + (size_t)lay * N;
    double* opt = d.opt + (size_t)lay * N;
    const double* newp = d.rhs + (size_t)lay * d.Np;
    const int iter = d.iter_nb[lay];
    double eq_error = __longlong_as_double((long long)d.eq_bits[lay]);
    double best = d.best_eq[lay];
    __syncthreads();
    // best-residual bookkeeping, iteration_control.h:105-109
    const bool better = eq_error < best;
    if (better) {
        best = eq_error;
        for (int k = threadIdx.x; k < N; k += CTL_THREADS) opt[k] = pop[k];
    }
    // residual + convergence max-reduction, :111-119; residual_list.push_front, :120
    const int hr = (d.head_res[lay] + RT_HIST - 1) % RT_HIST;
    double* res = ring_slot(d.hist_res, lay, N, hr);
    double pe = 0., re = 0.;
    for (int k = threadIdx.x; k < N; k += CTL_THREADS) {
        const double r = newp[k] - pop[k];
        res[k] = r;
        pe = fmax(pe, fabs(r));
        re = fmax(re, fabs(r / (pop[k] + 1.e-99)));
    }
    block_max2(pe, re);
    // next vector, :128-135
    const bool last = !(iter < max_iter - 1);
    __syncthreads();
    if (!last) {
        for (int k = threadIdx.x; k < N; k += CTL_THREADS) pop[k] = newp[k];
    }
    else {
        for (int k = threadIdx.x; k < N; k += CTL_THREADS) pop[k] = opt[k];
        eq_error = best;
    }
    const bool found = re < min_err;
    const bool stop = found || (iter + 1 >= max_iter);
    __syncthreads();
    if (threadIdx.x == 0) {
        d.head_res[lay] = hr;
        if (d.n_hist_res[lay] < RT_HIST) d.n_hist_res[lay]++;
        d.best_eq[lay] = best;
        d.iter_nb[lay] = iter + 1;
        d.eq_err[lay] = eq_error; d.pop_err[lay] = pe; d.rel_err[lay] = re;
        d.eq_bits[lay] = 0ull;
        if (stop) { d.active_flag[lay] = 0; d.conv[lay] = found ? 1 : 0; }
    }
    __syncthreads();
    if (!stop) pre_step(d, lay, accel);
    return stop;
}

