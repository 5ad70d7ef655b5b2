// misc_kernels.cu -- single-line intensity probe (rt_intensity) and the post-solve maser pass (rt_maser_postpass).
#include "physics.cuh"

// == virtual intensity_calc(up, low, level_pop, J) of the active scheme for one layer (iteration_lvg.cpp:163-185,
// overlap single-line overload :503-526).  which: 0 = single line, 1 = first line of a pair, 2 = second line.
__global__ void k_intensity_single(DevModel d, int lay, int u1, int l1, double A1ul, double A1lu, int u2, int l2,
                                   double A2ul, double A2lu, int which, const double* __restrict__ pop, double* out)
{
    const LayerCtx c = dev_layer_ctx(d, lay);
    double sigma[RT_MAX_DUST];
    // dust cross-section at the energy of line 1, evaluated here from the per-line table when the line is known,
    // else passed through out[1..ncomp] by the host
    for (int k = 0; k < d.ncomp; k++) sigma[k] = out[1 + k];
    const double e1 = d.energy[u1] - d.energy[l1];
    // one-shot evaluation: the reference's operation order exactly (physics.cuh, dev_intensity_single `exact`)
    if (which == 0) out[0] = dev_intensity_single(d, c, e1, A1ul, A1lu, pop[u1], pop[l1], sigma, true);
    else {
        double J1, J2;
        const double dxn = (d.energy[u1] - d.energy[l1] - d.energy[u2] + d.energy[l2]) * d.prm.speed_of_light;
        dev_intensity_pair(d, c, e1, dxn, A1ul, A1lu, A2ul, A2lu, pop[u1], pop[l1], pop[u2], pop[l2], sigma, J1, J2, true);
        out[0] = which == 1 ? J1 : J2;
    }
}

void launch_intensity_single(const DevModel& d, cudaStream_t s, int lay, int up, int low, double Aul, double Alu,
                             int pair_up, int pair_low, double pAul, double pAlu, int which, const double* pop_dev,
                             double* out_dev)
{
    g_rt_launches++;
    k_intensity_single<<<1, 1, 0, s>>>(d, lay, up, low, Aul, Alu, pair_up, pair_low, pAul, pAlu, which, pop_dev, out_dev);
}

// ------------------------------------------------------------------------------------------------------------
// K5: post-solve maser quantities of listed transitions, transition_data.cpp:210-378
// (calc_inv :210-223, calc_gain :240-294 incl. the oH2O 22 GHz special width :249-260, calc_line_profile :296-378).
// One CTA per transition.  Layer sums run in ascending layer order like the reference's loops.
// ------------------------------------------------------------------------------------------------------------
struct PostArgs {
    const double* pop;  // [L][N]
    const int *up, *low;
    const double* sigma; // [T][ncomp] dust cross-section at the transition energy
    const double *dz, *vel_n, *vgrid, *vel_turb; // [L], [L], [nb_freq], [L]
    double height, delta_aspect, gw_offset;
    int nb_freq, nb_aspect, gw_up, gw_low; // rt_set_gain_width_line: the line with the widened thermal width in calc_gain
    double *inv_arr, *gain_arr, *scalars, *tau_aspect, *tau_freq;
};

__global__ void __launch_bounds__(320) k_postpass(DevModel d, PostArgs a)
{
    const int t = blockIdx.x, L = d.L, N = d.N;
    extern __shared__ double sm[]; // line_opacity[L], dust_opacity[L], inv[L], gain[L]
    double* s_lo = sm; double* s_do = sm + L; double* s_inv = sm + 2 * L; double* s_gain = sm + 3 * L;
    __shared__ unsigned long long s_tmax[64];
    const int u = a.up[t], l = a.low[t];
    const double energy = d.energy[u] - d.energy[l]; // transition::energy, spectroscopy.cpp:1224-1238
    const double energy_th = energy * energy * energy;
    const double gu = d.g[u], gl = d.g[l];
    const double Aul = d.einst[(size_t)u * N + l];
    for (int lay = threadIdx.x; lay < L; lay += blockDim.x) {
        const double up_pop = a.pop[(size_t)lay * N + u], low_pop = a.pop[(size_t)lay * N + l];
        const double inv = up_pop / gu - low_pop / gl; // :218
        const double vel = d.vel_width[lay];           // pow(2kT/m + vt^2, 0.5), :261-262 == iteration_lvg.cpp:65
        double vel_gain = vel;
        if (u == a.gw_up && l == a.gw_low) {           // :257-260: thermal width + hyperfine offset, then the turbulent part
            const double th = sqrt(2. * d.prm.boltzmann_constant * d.temp_n[lay] / d.mol_mass) + a.gw_offset;
            vel_gain = sqrt(th * th + a.vel_turb[lay] * a.vel_turb[lay]);
        }
        const double line_gain = inv * gu * Aul * d.prm.onedivby_sqrt_pi * d.mol_conc[lay] / (energy_th * d.prm.eight_pi * vel_gain); // :266-268
        double d_abs = 0.;
        for (int c = 0; c < d.ncomp; c++) d_abs += a.sigma[(size_t)t * d.ncomp + c] * d.dust_conc[(size_t)lay * d.ncomp + c];
        s_inv[lay] = inv;
        s_gain[lay] = line_gain - d_abs;
        s_lo[lay] = inv * gu * Aul * d.mol_conc[lay] * d.prm.onedivby_sqrt_pi / (energy_th * d.prm.eight_pi * vel); // :335-337
        s_do[lay] = d_abs;
        if (a.inv_arr) a.inv_arr[(size_t)t * L + lay] = inv;
        if (a.gain_arr) a.gain_arr[(size_t)t * L + lay] = line_gain - d_abs;
    }
    if (threadIdx.x < 64) s_tmax[threadIdx.x] = 0ull;
    __syncthreads();
    if (threadIdx.x == 0 && a.scalars) {
        double inv = 0., gain = 0., tau_eff = 0., g = 0.;
        int hg = 0;
        for (int lay = 0; lay < L; lay++) {
            inv += s_inv[lay] * a.dz[lay];
            gain += s_gain[lay] * a.dz[lay];
            if (s_gain[lay] > 0.) tau_eff += s_gain[lay] * a.dz[lay];
            if (s_gain[lay] > g) { g = s_gain[lay]; hg = lay; }
        }
        if (g < 1.e-99) hg = 0;
        double* s = a.scalars + 5 * (size_t)t;
        s[0] = inv / a.height; s[1] = gain / a.height; s[2] = tau_eff; s[4] = hg;
    }
    const int n = threadIdx.x;
    if (n < a.nb_freq) {
        const double vel = a.vgrid[n];
        for (int i = 0; i < a.nb_aspect; i++) {
            const double aspect = 1. + a.delta_aspect * i;
            double tau = 0.;
            for (int lay = 0; lay < L; lay++) {
                const double x = (vel - a.vel_n[lay] / aspect) / d.vel_width[lay];
                const double profile = exp(-x * x);
                const double k = s_lo[lay] * profile - s_do[lay];
                if (k > 0.) tau += k * a.dz[lay] * aspect;
            }
            atomicMax(&s_tmax[i], (unsigned long long)__double_as_longlong(tau)); // tau >= 0
            if (i == 0 && a.tau_freq) a.tau_freq[(size_t)t * a.nb_freq + n] = tau;
        }
    }
    __syncthreads();
    if (threadIdx.x < a.nb_aspect) {
        const double v = __longlong_as_double((long long)s_tmax[threadIdx.x]);
        if (a.tau_aspect) a.tau_aspect[(size_t)t * a.nb_aspect + threadIdx.x] = v;
        if (threadIdx.x == 0 && a.scalars) a.scalars[5 * (size_t)t + 3] = v; // tau_max = tau_vs_aspect_ratio[0]
    }
}

void launch_postpass_impl(const DevModel& d, cudaStream_t s, const PostArgs& a, int nb_trans)
{
    if (nb_trans > 0) { g_rt_launches++; k_postpass<<<nb_trans, 320, 4 * d.L * sizeof(double), s>>>(d, a); }
}

// ------------------------------------------------------------------------------------------------------------
// FP64 ceilings (rt_device_fp64_peak): denominators of the LU roofline, measured in the same run as the bench.
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_peak_dfma(double* out, int iters)
{
    double a[8], x = 1.0000001 + threadIdx.x * 1e-9, y = 1e-9;
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = i + threadIdx.x;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) a[i] = fma(a[i], x, y);
    }
    double s = 0.;
#pragma unroll
    for (int i = 0; i < 8; i++) s += a[i];
    if (s == 12345.678) out[0] = s;
}

__global__ void __launch_bounds__(256) k_peak_dmma(double* out, int iters)
{
    double c[8][2];
#pragma unroll
    for (int i = 0; i < 8; i++) c[i][0] = c[i][1] = 0.;
    double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
    }
    double s = 0.;
#pragma unroll
    for (int i = 0; i < 8; i++) s += c[i][0] + c[i][1];
    if (s == 12345.678) out[0] = s;
}

extern "C" rt_status rt_device_fp64_peak(int device, int reps, double* dfma_tflops, double* dmma_tflops)
{
    if (cudaSetDevice(device) != cudaSuccess) return RT_ERR_CUDA;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return RT_ERR_CUDA;
    double* out = nullptr;
    if (cudaMalloc(&out, 64) != cudaSuccess) return RT_ERR_CUDA;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 4096;
    double best[2] = {0., 0.};
    for (int which = 0; which < 2; which++)
        for (int r = 0; r < reps + 1; r++) {
            cudaEventRecord(e0);
            if (which == 0) k_peak_dfma<<<blocks, threads>>>(out, iters);
            else k_peak_dmma<<<blocks, threads>>>(out, iters);
            cudaEventRecord(e1);
            if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(out); return RT_ERR_CUDA; }
            float ms = 0.f;
            cudaEventElapsedTime(&ms, e0, e1);
            // DFMA: 8 fma = 16 flop per thread-iteration; DMMA m8n8k4: 2*8*8*4 = 512 flop per warp instruction, 8 per iteration
            const double flop = which == 0 ? 16. * iters * (double)blocks * threads : 512. * 8. * iters * (double)blocks * (threads / 32);
            if (r > 0) best[which] = fmax(best[which], flop / (ms * 1e-3) * 1e-12);
        }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    if (dfma_tflops) *dfma_tflops = best[0];
    if (dmma_tflops) *dmma_tflops = best[1];
    return cudaGetLastError() == cudaSuccess ? RT_OK : RT_ERR_CUDA;
}
