// fused_kernels.cu -- the whole iteration of one layer in ONE persistent CTA, for molecules whose rate matrix fits in
// shared memory (Np <= 160: OH hyperfine N = 56, NH3/H2CO, H2O N = 150).
//
// Configs 1 and 3 of BASELINE.json (OH, N = 56, un-accelerated, up to 15000 iterations per layer) are launch-latency
// problems for the multi-kernel path: a step is ~10 launches of a few microseconds of work each.  Here one CTA per
// layer runs  intensity -> assembly + residual -> LU solve -> controller step  in a loop until its layer converges
// or hits the iteration cap, with the matrix, the populations and the radiative rates resident in shared memory and
// the controller state (history rings, Ng acceleration) in global memory exactly as in controller_kernels.cu.
// Layers are independent, so no CTA ever waits for another.
//
// Reference rows: a6-a9 + a16-a17 (coupling, assembly, residual), a10-a11 (LU), a12-a14.
#include "controller.cuh"
#include "physics.cuh"

#define FUSED_THREADS CTL_THREADS // the controller device code is written for CTL_THREADS cooperating threads

// shared-memory carve-up (doubles): M [Np][ldm] with ldm = Np + 1 (column Np = rhs; odd stride: conflict-free column
// walks), pop [Np], yd [R], yu [R]
__host__ __device__ inline size_t fused_smem_doubles(int Np, int R, bool with_c)
{
    return (size_t)Np * (Np + 1) + Np + 2 * (size_t)(R > 0 ? R : 1) + 64 + (with_c ? (size_t)Np * Np : 0);
}
__host__ __device__ inline bool fused_c_in_smem(int Np) { return Np <= 96; } // 2 x 73 KB at Np = 96

// Arg-max of the 64-bit keys of warp 0 (two REDUX), interchange of positions k <-> p in s_row, publication of the pivot's
// physical row, the displaced row, p and 1/pivot for parity k & 1.  key = |a| bits with the low byte replaced by
// 0xFF - (pos - k): larger value wins, ties (top 56 bits equal) -> smallest position, like the scalar loop.
struct FusedPiv { int phys_p, phys_old, p; double rp; };
__device__ __forceinline__ void pivot_finish(unsigned long long key, const double* M, int ldm, int k, short* s_row, FusedPiv* out, int lane)
{
    const unsigned hi = (unsigned)(key >> 32);
    const unsigned mhi = __reduce_max_sync(0xffffffffu, hi);
    const unsigned lo = hi == mhi ? (unsigned)key : 0u;
    const unsigned mlo = __reduce_max_sync(0xffffffffu, lo);
    if (lane == 0) {
        const int p = k + 0xFF - (int)(mlo & 0xFFu);
        // s_row itself is updated one step later by the caller (the updating warps are still reading it)
        const int phys_k = s_row[k], phys_p = s_row[p];
        const double piv = M[(size_t)phys_p * ldm + k];
        out->phys_p = phys_p; out->phys_old = phys_k; out->p = p;
        out->rp = piv != 0. ? __drcp_rn(piv) : 0.; // 0 encodes a zero pivot: the column is left as is
    }
}
__device__ __forceinline__ unsigned long long pivot_key(double v, int off)
{
    return ((unsigned long long)__double_as_longlong(fabs(v)) & ~0xFFull) | (unsigned long long)(0xFF - off);
}

// back substitution by one warp, NQ = Np / 32 position slots per lane, right-looking, no barriers
template <int NQ>
__device__ __forceinline__ void fused_backsolve(const double* M, int ldm, int Np, const short* rowf, double* gnew, int lane)
{
    double b[NQ], rd[NQ];
    const double* rowp[NQ];
#pragma unroll
    for (int q = 0; q < NQ; q++) {
        const int r = lane + 32 * q;
        rowp[q] = M + (size_t)rowf[r] * ldm;
        b[q] = rowp[q][Np];
        rd[q] = __drcp_rn(rowp[q][r]);
    }
#pragma unroll
    for (int qi = NQ - 1; qi >= 0; qi--) {
        for (int ii = 31; ii >= 0; ii--) {
            const int i = 32 * qi + ii;
            const double xi = __shfl_sync(0xffffffffu, b[qi] * rd[qi], ii);
#pragma unroll
            for (int q = 0; q < qi; q++) b[q] -= rowp[q][i] * xi;
            if (lane < ii) b[qi] -= rowp[qi][i] * xi;
            else if (lane == ii) b[qi] = xi;
        }
    }
#pragma unroll
    for (int q = 0; q < NQ; q++) gnew[lane + 32 * q] = b[q];
}

__global__ void __launch_bounds__(FUSED_THREADS) k_fused_solve(DevModel d, int nlay, int accel, int max_iter, double min_err)
{
    const int a = blockIdx.x;
    if (a >= nlay || a >= *d.n_active) return;
    const int lay = d.active[a];
    const int N = d.N, Np = d.Np, ldm = Np + 1, R = d.R;
    extern __shared__ __align__(16) double sm[];
    double* M = sm;                      // [Np][ldm]
    double* s_pop = M + (size_t)Np * ldm; // [Np]
    double* s_yd = s_pop + Np;           // [R]
    double* s_yu = s_yd + (R > 0 ? R : 1);
    __shared__ short s_row[160];  // position -> physical row of the LU
    __shared__ FusedPiv s_piv[2]; // pivot of the column, double-buffered by column parity
    __shared__ double s_eq[FUSED_THREADS / 32];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    // the collisional matrix of the layer is constant over the iteration: staged in shared memory when it fits
    const double* C = d.Cmat + (size_t)lay * Np * Np;
    if (fused_c_in_smem(Np)) {
        double* s_C = s_yu + (R > 0 ? R : 1);
        for (int e = tid; e < Np * Np; e += FUSED_THREADS) s_C[e] = C[e];
        C = s_C;
    }
    double* gpop = d.pop + (size_t)lay * N;
    double* gnew = d.rhs + (size_t)lay * Np;
    const LayerCtx ctx = dev_layer_ctx(d, lay);

    for (;;) {
        // ---- populations of this step (the controller left them in global memory) ----------------------------------
        for (int j = tid; j < Np; j += FUSED_THREADS) s_pop[j] = j < N ? gpop[j] : 0.;
        __syncthreads();
        // ---- K1a: mean intensities -> radiative rates (iteration_lvg.cpp:163-185, :428-501) ------------------------
        for (int gi = tid; gi < d.G; gi += FUSED_THREADS) {
            const int nb = d.grp[3 * gi], la = d.grp[3 * gi + 1];
            const int u1 = d.tr_u[la], l1 = d.tr_l[la];
            const double A1ul = d.tr_Aul[la], A1lu = d.tr_Alu[la];
            if (nb == 1) {
                const double J = dev_intensity_single(d, ctx, d.tr_E[la], A1ul, A1lu, s_pop[u1], s_pop[l1], d.tr_sigma + (size_t)la * d.ncomp);
                s_yd[la] = A1ul * (1. + J);
                s_yu[la] = A1lu * J;
            }
            else {
                const int lb = d.grp[3 * gi + 2];
                const int u2 = d.tr_u[lb], l2 = d.tr_l[lb];
                const double A2ul = d.tr_Aul[lb], A2lu = d.tr_Alu[lb];
                double J1, J2;
                dev_intensity_pair(d, ctx, d.tr_E[la], d.grp_dxn[gi], A1ul, A1lu, A2ul, A2lu, s_pop[u1], s_pop[l1], s_pop[u2], s_pop[l2],
                                   d.tr_sigma + (size_t)la * d.ncomp, J1, J2);
                s_yd[la] = A1ul * (1. + J1); s_yu[la] = A1lu * J1;
                s_yd[lb] = A2ul * (1. + J2); s_yu[lb] = A2lu * J2;
            }
        }
        __syncthreads();
        // ---- K1b: M = C + R, row 0 = 1, rhs = e0, df = e0 - M pop (iteration_lvg.cpp:112-161); one warp per row ------
        double eqmax = 0.;
        for (int row = warp; row < Np; row += FUSED_THREADS / 32) {
            double* Mrow = M + (size_t)row * ldm;
            const double* Crow = C + (size_t)row * Np;
            double dot = 0.;
            if (row == 0) {
                for (int j = lane; j < Np; j += 32) { const double v = j < N ? 1. : 0.; Mrow[j] = v; dot += v * s_pop[j]; }
            }
            else {
                for (int j = lane; j < Np; j += 32) { const double c = Crow[j]; Mrow[j] = c; dot += c * s_pop[j]; }
                if (row < N) {
                    __syncwarp();
                    double diag = 0.;
                    for (int e = d.row_ptr[row] + lane; e < d.row_ptr[row + 1]; e += 32) {
                        const int col = d.row_col[e], ln = d.row_line[e], line = ln >> 1;
                        const double off = (ln & 1) ? s_yd[line] : s_yu[line];
                        const double dg = (ln & 1) ? s_yu[line] : s_yd[line];
                        Mrow[col] = Crow[col] + off;
                        dot += off * s_pop[col];
                        diag += dg;
                    }
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) diag += __shfl_xor_sync(0xffffffffu, diag, o);
                    if (lane == 0) { Mrow[row] = Crow[row] - diag; dot -= diag * s_pop[row]; }
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
            if (lane == 0) {
                Mrow[Np] = row == 0 ? 1. : 0.;
                if (row < N) eqmax = fmax(eqmax, fabs((row == 0 ? 1. : 0.) - dot));
            }
        }
        if (lane == 0) s_eq[warp] = eqmax;
        __syncthreads();
        if (tid == 0) {
            double e = s_eq[0];
            for (int w = 1; w < FUSED_THREADS / 32; w++) e = fmax(e, s_eq[w]);
            d.eq_bits[lay] = (unsigned long long)__double_as_longlong(e);
        }
        // ---- K2: LU with row partial pivoting in shared memory, rhs eliminated on the fly (oracle/shim/linear_algebra.h).
        // Latency-optimised: ONE barrier per column.
        //  * rows are never moved: s_row[pos] = physical row at position pos;
        //  * warp 0 runs one column AHEAD: during step k it updates column k+1 itself, finds the pivot of column k+1
        //    (packed 64-bit keys, two REDUX max), interchanges the two positions and computes 1/pivot;
        //  * warps 1-7 apply step k to columns k+2 .. Np (rhs) as a 14 x 16 thread grid.
        // Values that agree in their top 56 bits count as ties (relative difference < 3e-14: immaterial for the pivot).
        {
            for (int r = tid; r < Np; r += FUSED_THREADS) s_row[r] = (short)r;
            __syncthreads();
            if (warp == 0) {
                unsigned long long key = 0ull;
                for (int pos = lane; pos < Np; pos += 32) { const unsigned long long kk = pivot_key(M[(size_t)pos * ldm], pos); key = kk > key ? kk : key; }
                pivot_finish(key, M, ldm, 0, s_row, &s_piv[0], lane);
            }
            __syncthreads();
            for (int k = 0; k < Np; k++) {
                const FusedPiv pv = s_piv[k & 1];
                const double rp = pv.rp;
                const double* Mp = M + (size_t)pv.phys_p * ldm;
                if (warp == 0) {
                    // interchange k <-> p of THIS column in s_row: the updating warps never read position k and
                    // override position p with phys_old, so nobody races with these two stores
                    if (lane == 0) { s_row[k] = (short)pv.phys_p; s_row[pv.p] = (short)pv.phys_old; }
                    __syncwarp();
                    if (k + 1 < Np) {
                        unsigned long long key = 0ull;
                        const double pk1 = Mp[k + 1];
                        for (int pos = k + 1 + lane; pos < Np; pos += 32) {
                            double* Mr = M + (size_t)(pos == pv.p ? pv.phys_old : s_row[pos]) * ldm;
                            const double l = Mr[k] * rp;
                            double v = Mr[k + 1];
                            if (l != 0.) { v -= l * pk1; Mr[k + 1] = v; }
                            const unsigned long long kk = pivot_key(v, pos - (k + 1));
                            key = kk > key ? kk : key;
                        }
                        __syncwarp();
                        pivot_finish(key, M, ldm, k + 1, s_row, &s_piv[(k + 1) & 1], lane);
                    }
                }
                else {
                    const int ty = (warp - 1) * 2 + (lane >> 4), tx = lane & 15;
                    for (int pos = k + 1 + ty; pos < Np; pos += 14) {
                        double* Mr = M + (size_t)(pos == pv.p ? pv.phys_old : s_row[pos]) * ldm;
                        const double l = Mr[k] * rp;
                        if (l != 0.)
                            for (int c = k + 2 + tx; c <= Np; c += 16) Mr[c] -= l * Mp[c];
                    }
                }
                __syncthreads();
            }
        }
        if (warp == 0) {
            switch (Np >> 5) {
            case 1: fused_backsolve<1>(M, ldm, Np, s_row, gnew, lane); break;
            case 2: fused_backsolve<2>(M, ldm, Np, s_row, gnew, lane); break;
            case 3: fused_backsolve<3>(M, ldm, Np, s_row, gnew, lane); break;
            case 4: fused_backsolve<4>(M, ldm, Np, s_row, gnew, lane); break;
            default: fused_backsolve<5>(M, ldm, Np, s_row, gnew, lane); break;
            }
        }
        __syncthreads();
        // ---- K3: controller step (iteration_control.h:84-137), shared with the multi-kernel path --------------------
        if (ctl_step_body(d, lay, accel, max_iter, min_err)) break;
        __syncthreads();
    }
}

void launch_fused_solve(const DevModel& d, cudaStream_t s, int nlay, int accel, int max_iter, double min_err)
{
    g_rt_launches++;
    k_fused_solve<<<nlay, FUSED_THREADS, fused_smem_doubles(d.Np, d.R, fused_c_in_smem(d.Np)) * sizeof(double), s>>>(d, nlay, accel, max_iter, min_err);
}

bool fused_supported(const DevModel& d)
{
    return d.Np <= 160 && fused_smem_doubles(d.Np, d.R, fused_c_in_smem(d.Np)) * sizeof(double) <= 220 * 1024;
}

cudaError_t fused_kernels_init()
{
    return cudaFuncSetAttribute(k_fused_solve, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
}
