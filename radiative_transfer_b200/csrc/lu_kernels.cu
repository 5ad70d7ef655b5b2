// lu_kernels.cu -- K2: batched FP64 dense solve  M x = e0, one matrix per active layer.
//
// Replaces lu_matrix_solve (missing from the reference; contract restated in oracle/shim/linear_algebra.h;
// call site iteration_lvg.cpp:97-101): LU with ROW PARTIAL PIVOTING (largest |a_ik| of the column, first
// occurrence wins), multipliers l = a_ik * (1/pivot), right-hand side eliminated on the fly, back substitution.
//
// Blocked right-looking algorithm on the padded [Np][Np] row-major matrices in HBM, panel width NB = 32:
//   for each panel k:  k_panel  : factor the (Np-32k) x 32 panel in shared memory (pivot search, row swaps,
//                                 rank-1 updates; the rhs rides along as a 33rd column), swap the trailing rows
//                      k_trsm   : U12 = L11^-1 A12                       (32 x m, unit lower triangular solve)
//                      k_gemm   : A22 -= L21 * U12                        (m x m x 32, the O(N^3) part)
//   k_backsolve: U x = y, blocked from the bottom.
// Work per layer: (2/3) Np^3 + O(Np^2) flop; HBM traffic ~ sum over panels of 16 m^2 bytes.
#include "rt_internal.h"
#include <cfloat>

#define PNL_THREADS 256
#define LU_PANEL_SMEM_MAX (226 * 1024) // 227 KB per CTA minus the kernel's static shared memory
#define PNL_LD 33 // 32 panel columns + rhs column; odd stride -> conflict-free column access

__global__ void __launch_bounds__(PNL_THREADS) k_panel(DevModel d, int k, int nlay_bound)
{
    const int a = blockIdx.x;
    if (a >= nlay_bound || a >= *d.n_active) return;
    const int lay = d.active[a];
    const int Np = d.Np, k0 = k * RT_NB, m = Np - k0;
    extern __shared__ double P[]; // [m][PNL_LD]
    __shared__ int s_piv[RT_NB];
    __shared__ double s_val[PNL_THREADS / 32];
    __shared__ int s_idx[PNL_THREADS / 32];
    double* __restrict__ A = d.Mmat + (size_t)lay * Np * Np;
    double* __restrict__ b = d.rhs + (size_t)lay * Np;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    for (int r = warp; r < m; r += PNL_THREADS / 32) P[r * PNL_LD + lane] = A[(size_t)(k0 + r) * Np + k0 + lane];
    for (int r = tid; r < m; r += PNL_THREADS) P[r * PNL_LD + 32] = b[k0 + r];
    __syncthreads();

    for (int j = 0; j < RT_NB; j++) {
        // ---- pivot search: max |P[r][j]|, r >= j; ties -> smallest r (first occurrence, as the scalar loop)
        double best = -1.;
        int bidx = 0x7fffffff;
        for (int r = j + tid; r < m; r += PNL_THREADS) {
            const double v = fabs(P[r * PNL_LD + j]);
            if (v > best) { best = v; bidx = r; } // r ascending per thread: strict '>' keeps the first
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bidx, o);
            if (ov > best || (ov == best && oi < bidx)) { best = ov; bidx = oi; }
        }
        if (lane == 0) { s_val[warp] = best; s_idx[warp] = bidx; }
        __syncthreads();
        if (warp == 0) {
            best = lane < PNL_THREADS / 32 ? s_val[lane] : -1.;
            bidx = lane < PNL_THREADS / 32 ? s_idx[lane] : 0x7fffffff;
#pragma unroll
            for (int o = 4; o > 0; o >>= 1) {
                const double ov = __shfl_xor_sync(0xffffffffu, best, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bidx, o);
                if (ov > best || (ov == best && oi < bidx)) { best = ov; bidx = oi; }
            }
            if (lane == 0) s_piv[j] = (bidx == 0x7fffffff) ? j : bidx; // all-NaN column: keep the diagonal
        }
        __syncthreads();
        const int p = s_piv[j];
        if (p != j && tid < PNL_LD) {
            const double t = P[j * PNL_LD + tid];
            P[j * PNL_LD + tid] = P[p * PNL_LD + tid];
            P[p * PNL_LD + tid] = t;
        }
        __syncthreads();
        const double piv = P[j * PNL_LD + j];
        if (piv != 0.) {
            const double rp = 1. / piv;
            for (int r = j + 1 + tid; r < m; r += PNL_THREADS) {
                double* row = P + r * PNL_LD;
                const double l = row[j] * rp;
                if (l != 0.) {
                    row[j] = l;
                    const double* prow = P + j * PNL_LD;
                    for (int c = j + 1; c <= RT_NB; c++) row[c] -= l * prow[c];
                }
            }
        }
        __syncthreads();
    }

    for (int r = warp; r < m; r += PNL_THREADS / 32) A[(size_t)(k0 + r) * Np + k0 + lane] = P[r * PNL_LD + lane];
    for (int r = tid; r < m; r += PNL_THREADS) b[k0 + r] = P[r * PNL_LD + 32];
    if (tid < RT_NB) d.ipiv[(size_t)lay * Np + k0 + tid] = k0 + s_piv[tid];
    // row interchanges for the columns to the right of the panel: columns are independent, so each thread
    // replays the 32 swaps in order for its own columns (coalesced across the warp for every row touched)
    for (int c = k0 + RT_NB + tid; c < Np; c += PNL_THREADS) {
        for (int j = 0; j < RT_NB; j++) {
            const int p = s_piv[j];
            if (p != j) {
                double* x = A + (size_t)(k0 + j) * Np + c;
                double* y = A + (size_t)(k0 + p) * Np + c;
                const double t = *x;
                *x = *y;
                *y = t;
            }
        }
    }
}

// U12 = L11^-1 * A12: grid (column tiles of 128, layers).  L11 is the unit-lower 32x32 block the panel kernel left
// in place.  Each thread owns one column of A12: loads it (coalesced across the CTA for every row), forward
// substitution entirely in registers, stores it back.
#define TRSM_THREADS 128
__global__ void __launch_bounds__(TRSM_THREADS) k_trsm(DevModel d, int k, int nlay_bound)
{
    const int a = blockIdx.y;
    if (a >= nlay_bound || a >= *d.n_active) return;
    const int lay = d.active[a];
    const int Np = d.Np, k0 = k * RT_NB;
    __shared__ double Ls[32][33];
    double* __restrict__ A = d.Mmat + (size_t)lay * Np * Np;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int r = warp; r < 32; r += TRSM_THREADS / 32) Ls[r][lane] = A[(size_t)(k0 + r) * Np + k0 + lane];
    __syncthreads();
    const int c = k0 + RT_NB + blockIdx.x * TRSM_THREADS + tid;
    if (c >= Np) return;
    double x[32];
#pragma unroll
    for (int r = 0; r < 32; r++) x[r] = A[(size_t)(k0 + r) * Np + c];
#pragma unroll
    for (int r = 1; r < 32; r++) {
        double s = x[r];
#pragma unroll
        for (int q = 0; q < r; q++) s -= Ls[r][q] * x[q];
        x[r] = s;
    }
#pragma unroll
    for (int r = 1; r < 32; r++) A[(size_t)(k0 + r) * Np + c] = x[r];
}

// A22 -= L21 * U12.  64 x 64 output tile per CTA, 256 threads, 4 x 4 register tile per thread, K = 32.
#define GT 64
__global__ void __launch_bounds__(256) k_gemm(DevModel d, int k, int nlay_bound)
{
    const int a = blockIdx.z;
    if (a >= nlay_bound || a >= *d.n_active) return;
    const int lay = d.active[a];
    const int Np = d.Np, k0 = k * RT_NB, t0 = k0 + RT_NB;
    const int i0 = t0 + blockIdx.y * GT, j0 = t0 + blockIdx.x * GT;
    __shared__ double As[RT_NB][GT + 1]; // As[q][i] = L21[i][q]  (transposed for conflict-free reads)
    __shared__ double Bs[RT_NB][GT];     // Bs[q][j] = U12[q][j]
    double* __restrict__ A = d.Mmat + (size_t)lay * Np * Np;
    const int tid = threadIdx.x;
    // load L21 tile: 64 rows x 32 cols
    for (int e = tid; e < GT * RT_NB; e += 256) {
        const int r = e >> 5, q = e & 31;
        As[q][r] = (i0 + r < Np) ? A[(size_t)(i0 + r) * Np + k0 + q] : 0.;
    }
    for (int e = tid; e < GT * RT_NB; e += 256) {
        const int q = e >> 6, c = e & 63;
        Bs[q][c] = (j0 + c < Np) ? A[(size_t)(k0 + q) * Np + j0 + c] : 0.;
    }
    __syncthreads();
    const int tx = tid & 15, ty = tid >> 4; // thread tile: rows ty*4.., cols tx*4..
    double acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = 0.;
#pragma unroll 8
    for (int q = 0; q < RT_NB; q++) {
        double av[4], bv[4];
#pragma unroll
        for (int i = 0; i < 4; i++) av[i] = As[q][ty * 4 + i];
#pragma unroll
        for (int j = 0; j < 4; j++) bv[j] = Bs[q][tx * 4 + j];
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
            for (int j = 0; j < 4; j++) acc[i][j] += av[i] * bv[j];
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int r = i0 + ty * 4 + i;
        if (r < Np) {
            double* row = A + (size_t)r * Np + j0 + tx * 4;
            if (j0 + tx * 4 + 3 < Np) {
                double2 v0 = *reinterpret_cast<double2*>(row), v1 = *reinterpret_cast<double2*>(row + 2);
                v0.x -= acc[i][0]; v0.y -= acc[i][1]; v1.x -= acc[i][2]; v1.y -= acc[i][3];
                *reinterpret_cast<double2*>(row) = v0;
                *reinterpret_cast<double2*>(row + 2) = v1;
            }
        }
    }
}

// U x = y, blocked from the bottom; one CTA per layer.  x overwrites rhs.
__global__ void __launch_bounds__(256) k_backsolve(DevModel d, int nlay_bound)
{
    const int a = blockIdx.x;
    if (a >= nlay_bound || a >= *d.n_active) return;
    const int lay = d.active[a];
    const int Np = d.Np, nb = Np / RT_NB;
    extern __shared__ double xs[]; // [Np]
    __shared__ double Us[32][33];
    __shared__ double ss[32];
    const double* __restrict__ A = d.Mmat + (size_t)lay * Np * Np;
    double* __restrict__ b = d.rhs + (size_t)lay * Np;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int kb = nb - 1; kb >= 0; kb--) {
        const int r0 = kb * RT_NB;
        // s[r] = y[r] - sum_{c >= r0+32} U[r][c] x[c]
        for (int rr = warp; rr < 32; rr += 8) {
            const double* row = A + (size_t)(r0 + rr) * Np;
            double s = 0.;
            for (int c = r0 + RT_NB + lane; c < Np; c += 32) s += row[c] * xs[c];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane == 0) ss[rr] = b[r0 + rr] - s;
            Us[rr][lane] = row[r0 + lane];
        }
        __syncthreads();
        if (warp == 0) {
            double s = ss[lane];
            double xv = 0.;
            for (int c = 31; c >= 0; c--) {
                const double sc = __shfl_sync(0xffffffffu, s, c);
                const double xc = sc / Us[c][c];
                if (lane == c) xv = xc;
                if (lane < c) s -= Us[lane][c] * xc;
            }
            xs[r0 + lane] = xv;
            b[r0 + lane] = xv;
        }
        __syncthreads();
    }
}

void launch_lu_solve(rt_model* m, cudaStream_t s, int nlay_bound)
{
    const DevModel& d = m->d;
    const int Np = d.Np, nb = Np / RT_NB;
    const int c_panel = rt_prof_class(m, "lu_panel"); 
    const int c_trsm = rt_prof_class(m, "lu_trsm");
    const int c_gemm = rt_prof_class(m, "lu_gemm"), c_back = rt_prof_class(m, "lu_backsolve");
    for (int k = 0; k < nb; k++) {
        const int mrows = Np - k * RT_NB;
        const int mt = mrows - RT_NB;
        const size_t smem = (size_t)mrows * PNL_LD * sizeof(double);
        rt_prof_begin(m, c_panel, s, 32. * 32. * mrows, 16. * 32. * mrows);
        g_rt_launches++;
        k_panel<<<nlay_bound, PNL_THREADS, smem, s>>>(d, k, nlay_bound);
        rt_prof_end(m, c_panel, s);
        if (mt > 0) {
            rt_prof_begin(m, c_trsm, s, 32. * 32. * mt, 16. * 32. * mt);
            g_rt_launches++;
            k_trsm<<<dim3((mt + TRSM_THREADS - 1) / TRSM_THREADS, nlay_bound), TRSM_THREADS, 0, s>>>(d, k, nlay_bound);
            rt_prof_end(m, c_trsm, s);
            const int nt = (mt + GT - 1) / GT;
            rt_prof_begin(m, c_gemm, s, 2. * 32. * mt * mt, 16. * (double)mt * mt);
            g_rt_launches++;
            k_gemm<<<dim3(nt, nt, nlay_bound), 256, 0, s>>>(d, k, nlay_bound);
            rt_prof_end(m, c_gemm, s);
        }
    }
    rt_prof_begin(m, c_back, s, (double)Np * Np, 4. * Np * Np);
    g_rt_launches++;
    k_backsolve<<<nlay_bound, 256, Np * sizeof(double), s>>>(d, nlay_bound);
    rt_prof_end(m, c_back, s);
}

// dynamic shared memory opt-in (up to 227 KB per CTA on sm_100)
int lu_max_np() { return LU_PANEL_SMEM_MAX / (PNL_LD * (int)sizeof(double)) / RT_NB * RT_NB; }

cudaError_t lu_kernels_init()
{
    return cudaFuncSetAttribute(k_panel, cudaFuncAttributeMaxDynamicSharedMemorySize, LU_PANEL_SMEM_MAX);
}
