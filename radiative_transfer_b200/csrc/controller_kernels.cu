// controller_kernels.cu -- K3: device-resident restatement of iteration_control<T> (iteration_control.h:69-242):
// history rings, best-residual bookkeeping, convergence max-reduction, Ng-type acceleration, active-layer list.
// One CTA per layer.  The iteration of layer l is independent of every other layer, so
// each layer carries its own iter_nb / nb_after_accel / converged flag and drops out of the active list alone.
//
// Per controller step the host enqueues:  k_intensity, k_assemble, LU kernels, k_ctl_step, k_compact.
// k_ctl_step does the tail of next_step for iteration n (iteration_control.h:105-136) and, for layers that go
// on, the head of next_step for iteration n+1 (:92-98: history push + acceleration), so that `pop` always holds
// the vector the next assembly must use.
#include "rt_internal.h"

#define CTL_THREADS 256

__device__ __forceinline__ double* ring_slot(double* base, int lay, int N, int slot)
{
    return base + ((size_t)lay * RT_HIST + slot) * N;
}

// max-reduce two non-negative values over the CTA; result valid in all threads
__device__ __forceinline__ void block_max2(double& a, double& b)
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
__device__ void solve4(double* A, double* b)
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
__device__ void accel_step(const DevModel& d, int lay, double* pop)
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
__device__ void pre_step(const DevModel& d, int lay, int accel)
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

// (re)start the controller of layers [lay_begin, lay_end): calculate_populations preamble, iteration_control.h:202-213
__global__ void __launch_bounds__(CTL_THREADS) k_ctl_init(DevModel d, int accel)
{
    const int lay = blockIdx.x;
    if (d.active_flag[lay] == 0) return; // only layers flagged by k_set_active_range / k_set_active_failed
    const int N = d.N;
    if (threadIdx.x == 0) {
        d.iter_nb[lay] = 0; d.nb_after[lay] = 0;
        d.head_pop[lay] = 0; d.head_res[lay] = 0; d.n_hist_pop[lay] = 0; d.n_hist_res[lay] = 0;
        d.best_eq[lay] = 1.; d.eq_err[lay] = 0.; d.pop_err[lay] = 0.; d.rel_err[lay] = 0.;
        d.conv[lay] = 0; d.active_flag[lay] = 1; d.eq_bits[lay] = 0ull;
    }
    double* opt = d.opt + (size_t)lay * N;
    for (int k = threadIdx.x; k < N; k += CTL_THREADS) opt[k] = 0.;
    __syncthreads();
    pre_step(d, lay, accel);
}

__global__ void __launch_bounds__(CTL_THREADS) k_ctl_step(DevModel d, int nlay_bound, int accel, int max_iter, double min_err)
{
    const int a = blockIdx.x;
    if (a >= nlay_bound || a >= *d.n_active) return;
    const int lay = d.active[a];
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
}

// Ordered compaction of active_flag[0..L) into active[] / n_active (single CTA; L is a few hundred).
__global__ void __launch_bounds__(1024) k_compact(DevModel d)
{
    __shared__ int s_scan[1024];
    __shared__ int s_base;
    if (threadIdx.x == 0) s_base = 0;
    __syncthreads();
    for (int c0 = 0; c0 < d.L; c0 += 1024) {
        const int lay = c0 + threadIdx.x;
        const int f = (lay < d.L) ? d.active_flag[lay] : 0;
        s_scan[threadIdx.x] = f;
        __syncthreads();
        for (int o = 1; o < 1024; o <<= 1) {
            const int v = threadIdx.x >= o ? s_scan[threadIdx.x - o] : 0;
            __syncthreads();
            s_scan[threadIdx.x] += v;
            __syncthreads();
        }
        if (f) d.active[s_base + s_scan[threadIdx.x] - 1] = lay;
        __syncthreads();
        if (threadIdx.x == 1023) s_base += s_scan[1023];
        __syncthreads();
    }
    if (threadIdx.x == 0) *d.n_active = s_base;
}

// mark layers [lay_begin, lay_end) active, all others inactive (used by rt_step / rt_assemble / chained mode)
__global__ void k_set_active_range(DevModel d, int lay_begin, int lay_end)
{
    const int lay = blockIdx.x * blockDim.x + threadIdx.x;
    if (lay < d.L) d.active_flag[lay] = (lay >= lay_begin && lay < lay_end) ? 1 : 0;
}

// keep only layers that did not converge (retry policy, radiative_transfer.cpp:259-276)
__global__ void k_set_active_failed(DevModel d, int lay_begin, int lay_end)
{
    const int lay = blockIdx.x * blockDim.x + threadIdx.x;
    if (lay < d.L) d.active_flag[lay] = (lay >= lay_begin && lay < lay_end && d.conv[lay] == 0) ? 1 : 0;
}

// boltzmann_populations, spectroscopy.cpp:1245-1259 (sequential partition sum, as the reference)
__global__ void k_boltzmann(DevModel d)
{
    const int lay = blockIdx.x;
    if (d.active_flag[lay] == 0) return;
    double* pop = d.pop + (size_t)lay * d.N;
    __shared__ double s_sum;
    const double t = d.temp_n[lay];
    for (int i = threadIdx.x; i < d.N; i += blockDim.x) pop[i] = d.g[i] * exp(-d.energy[i] * d.prm.cm_inverse_to_kelvins / t);
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.;
        for (int i = 0; i < d.N; i++) s = __dadd_rn(s, pop[i]);
        s_sum = s;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < d.N; i += blockDim.x) pop[i] = pop[i] / s_sum;
}

__global__ void k_set_rhs(DevModel d, int nlay_bound)
{
    const int a = blockIdx.y;
    if (a >= nlay_bound || a >= *d.n_active) return;
    const int lay = d.active[a];
    // b = e0 rides as column Np of the matrix; columns Np+1 .. ld-1 are zero padding
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < d.Np) {
        double* row = d.Mmat + ((size_t)lay * d.Np + i) * d.ld + d.Np;
        row[0] = (i == 0) ? 1. : 0.;
        for (int c = 1; c < d.ld - d.Np; c++) row[c] = 0.;
    }
}

// matrix[0][i] = 1 (boundary_layer_populations, iteration_control.cpp:81-83); padded columns stay 0
__global__ void k_set_row0(DevModel d, int nlay_bound)
{
    const int a = blockIdx.y;
    if (a >= nlay_bound || a >= *d.n_active) return;
    const int lay = d.active[a];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < d.Np) d.Mmat[(size_t)lay * d.Np * d.ld + i] = (i < d.N) ? 1. : 0.;
}

__global__ void k_copy_rhs_to_pop(DevModel d, int nlay_bound)
{
    const int a = blockIdx.y;
    if (a >= nlay_bound || a >= *d.n_active) return;
    const int lay = d.active[a];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < d.N) d.pop[(size_t)lay * d.N + i] = d.rhs[(size_t)lay * d.Np + i];
}

void launch_ctl_init_flagged(const DevModel& d, cudaStream_t s, int accel) { g_rt_launches++; k_ctl_init<<<d.L, CTL_THREADS, 0, s>>>(d, accel); }
void launch_ctl_step(const DevModel& d, cudaStream_t s, int nlay_bound, int accel, int max_iter, double min_err)
{
    g_rt_launches++;
    k_ctl_step<<<nlay_bound, CTL_THREADS, 0, s>>>(d, nlay_bound, accel, max_iter, min_err);
}
void launch_compact(const DevModel& d, cudaStream_t s) { g_rt_launches++; k_compact<<<1, 1024, 0, s>>>(d); }
void launch_set_active_range(const DevModel& d, cudaStream_t s, int b, int e) { g_rt_launches++; k_set_active_range<<<(d.L + 255) / 256, 256, 0, s>>>(d, b, e); }
void launch_set_active_failed(const DevModel& d, cudaStream_t s, int b, int e) { g_rt_launches++; k_set_active_failed<<<(d.L + 255) / 256, 256, 0, s>>>(d, b, e); }
void launch_boltzmann(const DevModel& d, cudaStream_t s) { g_rt_launches++; k_boltzmann<<<d.L, 256, 0, s>>>(d); }
void launch_set_rhs(const DevModel& d, cudaStream_t s, int nlay_bound)
{
    g_rt_launches++;
    k_set_rhs<<<dim3((d.Np + 255) / 256, nlay_bound), 256, 0, s>>>(d, nlay_bound);
}
void launch_set_row0(const DevModel& d, cudaStream_t s, int nlay_bound)
{
    g_rt_launches++;
    k_set_row0<<<dim3((d.Np + 255) / 256, nlay_bound), 256, 0, s>>>(d, nlay_bound);
}
void launch_copy_rhs_to_pop(const DevModel& d, cudaStream_t s, int nlay_bound)
{
    g_rt_launches++;
    k_copy_rhs_to_pop<<<dim3((d.N + 255) / 256, nlay_bound), 256, 0, s>>>(d, nlay_bound);
}
