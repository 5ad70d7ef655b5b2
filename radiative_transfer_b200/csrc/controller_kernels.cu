// controller_kernels.cu -- K3: device-resident restatement of iteration_control<T> (iteration_control.h:69-242):
// history rings, best-residual bookkeeping, convergence max-reduction, Ng-type acceleration, active-layer list.
// One CTA per layer.  The iteration of layer l is independent of every other layer, so
// each layer carries its own iter_nb / nb_after_accel / converged flag and drops out of the active list alone.
//
// Per controller step the host enqueues:  k_intensity, k_assemble, LU kernels, k_ctl_step, k_compact.
// k_ctl_step does the tail of next_step for iteration n (iteration_control.h:105-136) and, for layers that go
// on, the head of next_step for iteration n+1 (:92-98: history push + acceleration), so that `pop` always holds
// the vector the next assembly must use.
#include "controller.cuh"

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
    rt_pdl_sync();
    const int a = blockIdx.x;
    if (a >= nlay_bound || a >= *d.n_active) return;
    ctl_step_body(d, d.active[a], accel, max_iter, min_err);
}

// Ordered compaction of active_flag[0..L) into active[] / n_active (single CTA; L is a few hundred).
__global__ void __launch_bounds__(1024) k_compact(DevModel d)
{
    rt_pdl_sync();
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

// the layers listed in d.lay_list[0..n) active, all others inactive (chain step l of every stacked model)
__global__ void k_set_active_list(DevModel d, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) d.active_flag[d.lay_list[i]] = 1;
}

// warm start of the chain, radiative_transfer.cpp:247-249: pop[lay] <- pop[lay - 1] for the listed layers
__global__ void k_copy_prev_pop(DevModel d, int n)
{
    const int i = blockIdx.y;
    if (i >= n) return;
    const int lay = d.lay_list[i];
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < d.N && lay > 0) d.pop[(size_t)lay * d.N + k] = d.pop[(size_t)(lay - 1) * d.N + k];
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
    rt_launch(k_ctl_step, nlay_bound, CTL_THREADS, 0, s, d, nlay_bound, accel, max_iter, min_err);
}
void launch_compact(const DevModel& d, cudaStream_t s) { g_rt_launches++; rt_launch(k_compact, 1, 1024, 0, s, d); }
void launch_set_active_range(const DevModel& d, cudaStream_t s, int b, int e) { g_rt_launches++; k_set_active_range<<<(d.L + 255) / 256, 256, 0, s>>>(d, b, e); }
void launch_set_active_failed(const DevModel& d, cudaStream_t s, int b, int e) { g_rt_launches++; k_set_active_failed<<<(d.L + 255) / 256, 256, 0, s>>>(d, b, e); }
void launch_set_active_list(const DevModel& d, cudaStream_t s, int n)
{
    launch_set_active_range(d, s, 0, 0);
    if (n <= 0) return;
    g_rt_launches++;
    k_set_active_list<<<(n + 255) / 256, 256, 0, s>>>(d, n);
}
void launch_copy_prev_pop(const DevModel& d, cudaStream_t s, int n)
{
    if (n <= 0) return;
    g_rt_launches++;
    k_copy_prev_pop<<<dim3((d.N + 255) / 256, n), 256, 0, s>>>(d, n);
}
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
