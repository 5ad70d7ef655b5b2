// coupling_kernels.cu -- K0 (collisional matrix, once per cloud) and K1 (per-iteration coupling pass:
// mean intensities of all lines, rate-matrix assembly M = C + R(pop), residual df = e0 - M*pop).
//
// Reference rows covered: a3-a5 (set_gas_param / get_rate / mixing rules) in
// k_coll_offdiag + k_coll_diag; a6-a8 (intensity_calc / dust absorption / get_esc_func) in k_intensity;
// a9 (operator(): assembly + residual) in k_assemble; a16-a17 (OH hyperfine overlap) in k_intensity.
#include "physics.cuh"

// ------------------------------------------------------------------------------------------------------------
// K0: collisional rates.  One thread per (i > j) pair of one layer; j fastest so that the packed table index
// p = i(i-1)/2 + j and the row write C[i][j] are coalesced.
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ double tab_rate(const DevTable& t, int up, int low, int lo, double temp)
{
    // collision_data::get_rate(first, sec, lo, temp), coll_rates.cpp:54-59, temp clamped by the caller (:188)
    const size_t p = (size_t)up * (up - 1) / 2 + low;
    const double tt = temp < t.max_temp ? temp : t.max_temp;
    if (t.interp == RT_INTERP_CUBIC_SPLINE) {
        // collision_data_cub_spline::get_rate, coll_rates.cpp:121-132 (lo <= jmax - 2 by collision_data::locate)
        const double h = t.tgrid[lo + 1] - t.tgrid[lo];
        const double a = (t.tgrid[lo + 1] - tt) / h;
        const double b = 1. - a;
        const double c0 = t.coeff_t[(size_t)lo * t.imax + p], c1 = t.coeff_t[(size_t)(lo + 1) * t.imax + p];
        const double s0 = t.deriv_t[(size_t)lo * t.imax + p], s1 = t.deriv_t[(size_t)(lo + 1) * t.imax + p];
        return fabs(a * c0 + b * c1 + (1. / 6.) * ((a * a * a - a) * s0 + (b * b * b - b) * s1) * (h * h));
    }
    return t.coeff_t[(size_t)lo * t.imax + p] + t.deriv_t[(size_t)lo * t.imax + p] * (tt - t.tgrid[lo]);
}

__device__ __forceinline__ void rate_neutrals(const DevModel& d, int lay, int up, int low, double& down, double& upr)
{
    const double tn = d.temp_n[lay];
    const double* conc = d.conc + (size_t)lay * d.K;
    const int* idx = d.tidx + (size_t)lay * d.K;
    const DevTable* T = d.tab;
    double dn = 0.;
    switch (d.rule) {
    case RT_RULE_CH3OH: // coll_rates_ch3oh.cpp:484-533 with USE_TEMPER_EXTRAP_CH3OH == 0
        if (d.v[up] == d.v[low]) {
            if (d.v[up] == 0 && d.jq[up] <= 9 && d.jq[low] <= 9)
                dn = tab_rate(T[1], up, low, idx[1], tn) * conc[1] + tab_rate(T[2], up, low, idx[2], tn) * conc[2];
            else
                dn = tab_rate(T[1], up, low, idx[1], tn) * (conc[1] + conc[2]);
            dn += tab_rate(T[0], up, low, idx[0], tn) * conc[0];
        }
        else dn = tab_rate(T[0], up, low, idx[0], tn) * (conc[0] + conc[1] + 3. * conc[2]);
        break;
    case RT_RULE_OH_HF: // coll_rates_oh.cpp:392-407
        if (up < T[0].nb_lev) dn = tab_rate(T[0], up, low, idx[0], tn) * conc[0];
        if (up < T[1].nb_lev)
            dn += tab_rate(T[1], up, low, idx[1], tn) * conc[1] + tab_rate(T[2], up, low, idx[2], tn) * conc[2];
        break;
    case RT_RULE_H2O: // coll_rates_h2o.cpp:530-548
        if (up < 45)
            dn = tab_rate(T[0], up, low, idx[0], tn) * conc[0] + tab_rate(T[2], up, low, idx[2], tn) * conc[2]
                 + tab_rate(T[3], up, low, idx[3], tn) * conc[3] + tab_rate(T[5], up, low, idx[5], tn) * conc[5];
        else
            dn = tab_rate(T[1], up, low, idx[1], tn) * conc[1] + tab_rate(T[4], up, low, idx[4], tn) * conc[4];
        break;
    case RT_RULE_NH3: // coll_rates_nh3.cpp:273-290: the three hydrogen tables share the level limit of table 1
        if (up < T[0].nb_lev) dn = tab_rate(T[0], up, low, idx[0], tn) * conc[0];
        if (up < T[1].nb_lev)
            dn += tab_rate(T[1], up, low, idx[1], tn) * conc[1] + tab_rate(T[2], up, low, idx[2], tn) * conc[2]
                  + tab_rate(T[3], up, low, idx[3], tn) * conc[3];
        break;
    case RT_RULE_H2CO: // coll_rates_h2co.cpp:170-185
        if (up < T[0].nb_lev) dn = tab_rate(T[0], up, low, idx[0], tn) * conc[0];
        if (up < T[1].nb_lev)
            dn += tab_rate(T[1], up, low, idx[1], tn) * conc[1] + tab_rate(T[2], up, low, idx[2], tn) * conc[2];
        break;
    default: // coll_rates.cpp:181-197
        for (int k = 0; k < d.nb1; k++)
            if (up < T[k].nb_lev) dn += tab_rate(T[k], up, low, idx[k], tn) * conc[k];
    }
    if (dn > d.prm.min_collision_rate) {
        down = dn;
        upr = dn * exp((d.energy[low] - d.energy[up]) * d.prm.cm_inverse_to_kelvins / tn) * d.g[up] / ((double)d.g[low]);
    }
    else down = upr = 0.;
}

__device__ __forceinline__ void rate_electrons(const DevModel& d, int lay, int up, int low, double& down, double& upr)
{
    // coll_rates.cpp:199-217: first applicable electron table only
    const double te = d.temp_el[lay];
    const double* conc = d.conc + (size_t)lay * d.K;
    const int* idx = d.tidx + (size_t)lay * d.K;
    double dn = 0.;
    for (int k = d.nb1; k < d.nb2; k++)
        if (up < d.tab[k].nb_lev) {
            dn = tab_rate(d.tab[k], up, low, idx[k], te) * conc[k];
            break;
        }
    if (dn > d.prm.min_collision_rate) {
        down = dn;
        upr = dn * exp((d.energy[low] - d.energy[up]) * d.prm.cm_inverse_to_kelvins / te) * d.g[up] / ((double)d.g[low]);
    }
    else down = upr = 0.;
}

// grid (ceil(N/32), ceil(N/32), L), block (32, 8): tile (bi, bj) of the lower triangle; writes C[i][j] = up-rate
// j->i and, through a shared-memory transpose, C[j][i] = down-rate i->j, both as coalesced row segments.
// half_A != 0: boundary_layer_populations variant (iteration_control.cpp:72-78): neutrals only, + 0.5*A_ul.
// Layers: lay0 + blockIdx.z, or (bound > 0) the entries lay0 + blockIdx.z < min(bound, *n_active) of the compacted active list.
__device__ __forceinline__ int coll_layer(const DevModel& d, int idx, int bound)
{
    if (bound <= 0) return idx < d.L ? idx : -1;
    return (idx < bound && idx < *d.n_active) ? d.active[idx] : -1;
}

__global__ void __launch_bounds__(256) k_coll_offdiag(DevModel d, double* __restrict__ out, int ldo, int half_A,
                                                      const double* __restrict__ einst, int lay0, int bound)
{
    const int bi = blockIdx.y, bj = blockIdx.x;
    if (bj > bi) return;
    const int lay = coll_layer(d, lay0 + blockIdx.z, bound);
    if (lay < 0) return;
    __shared__ double s_down[32][33];
    double* C = out + (size_t)lay * d.Np * ldo;
    const int tx = threadIdx.x, ty = threadIdx.y;
    for (int r = ty; r < 32; r += 8) {
        const int i = bi * 32 + r, j = bj * 32 + tx;
        double down = 0., up = 0.;
        if (i < d.N && j < i) {
            rate_neutrals(d, lay, i, j, down, up);
            if (!half_A) {
                double a, b;
                rate_electrons(d, lay, i, j, a, b);
                down += a;
                up += b;
            }
            else down = 0.5 * einst[(size_t)i * d.N + j] + down;
            C[(size_t)i * ldo + j] = up;
        }
        s_down[r][tx] = down;
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int j = bj * 32 + r, i = bi * 32 + tx; // transposed: row j, column i
        if (i < d.N && j < i) C[(size_t)j * ldo + i] = s_down[tx][r];
    }
}

// Diagonal = -(sum of the other entries of the column) (iteration_lvg.cpp:127-131: every rate leaves its source
// level through the diagonal of the source column); padding block = identity.  One thread per column, coalesced.
__global__ void k_coll_diag(DevModel d, double* __restrict__ out, int ldo, int lay0, int bound)
{
    const int lay = coll_layer(d, lay0 + blockIdx.y, bound);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (lay < 0 || i >= d.Np) return;
    double* C = out + (size_t)lay * d.Np * ldo;
    if (i >= d.N) {
        for (int r = 0; r < d.Np; r++) C[(size_t)r * ldo + i] = (r == i) ? 1. : 0.;
        return;
    }
    double s = 0.;
    for (int r = 0; r < d.N; r++)
        if (r != i) s += C[(size_t)r * ldo + i];
    C[(size_t)i * ldo + i] = -s;
    for (int r = d.N; r < d.Np; r++) C[(size_t)r * ldo + i] = 0.;
}

// bound <= 0: every layer of the cloud; bound > 0: the first `bound` entries of the compacted active list (the start
// vectors of a chained solve or of the retry pass touch only the layers they are for).  Layers ride in gridDim.z / .y,
// which the hardware caps at 65535: larger (stacked) clouds go in chunks.
void launch_coll_matrix(const DevModel& d, cudaStream_t s, double* out, int half_A, int bound)
{
    const int ldo = (out == d.Mmat) ? d.ld : d.Np; // the boundary-start variant assembles straight into Mmat
    const int nt = (d.N + 31) / 32;
    const int nl = bound > 0 ? bound : d.L;
    for (int lay0 = 0; lay0 < nl; lay0 += 65535) {
        const int n = nl - lay0 < 65535 ? nl - lay0 : 65535;
        dim3 grid(nt, nt, n), block(32, 8);
        g_rt_launches++;
        k_coll_offdiag<<<grid, block, 0, s>>>(d, out, ldo, half_A, d.einst, lay0, bound);
        dim3 g2((d.Np + 127) / 128, n);
        g_rt_launches++;
        k_coll_diag<<<g2, 128, 0, s>>>(d, out, ldo, lay0, bound);
    }
}

// ------------------------------------------------------------------------------------------------------------
// K1a: mean intensity of every line of every active layer.  One thread per group (1 line, or 2 overlapping OH
// hyperfine lines); writes the two radiative rates of each line
//     yd = A_ul (1 + J)   (u -> l),   yu = A_lu J   (l -> u)          iteration_lvg.cpp:137-144
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void intensity_group(const DevModel& d, int lay, int gi, const double* __restrict__ pop, bool exact)
{
    const LayerCtx c = dev_layer_ctx(d, lay);
    const int nb = d.grp[3 * gi], la = d.grp[3 * gi + 1];
    const int u1 = d.tr_u[la], l1 = d.tr_l[la];
    const double A1ul = d.tr_Aul[la], A1lu = d.tr_Alu[la];
    double* yd = d.yd + (size_t)lay * d.R;
    double* yu = d.yu + (size_t)lay * d.R;
    if (nb == 1) {
        const double J = dev_intensity_single(d, c, d.tr_E[la], A1ul, A1lu, pop[u1], pop[l1], d.tr_sigma + (size_t)la * d.ncomp, exact);
        yd[la] = A1ul * (1. + J);
        yu[la] = A1lu * J;
    }
    else {
        const int lb = d.grp[3 * gi + 2];
        const int u2 = d.tr_u[lb], l2 = d.tr_l[lb];
        const double A2ul = d.tr_Aul[lb], A2lu = d.tr_Alu[lb];
        double J1, J2;
        dev_intensity_pair(d, c, d.tr_E[la], d.grp_dxn[gi], A1ul, A1lu, A2ul, A2lu, pop[u1], pop[l1], pop[u2], pop[l2],
                           d.tr_sigma + (size_t)la * d.ncomp, J1, J2, exact);
        yd[la] = A1ul * (1. + J1);
        yu[la] = A1lu * J1;
        yd[lb] = A2ul * (1. + J2);
        yu[lb] = A2lu * J2;
    }
}

__global__ void __launch_bounds__(128) k_intensity(DevModel d, int nlay_bound)
{
    rt_pdl_sync();
    const int a = blockIdx.y;
    if (a >= nlay_bound || a >= *d.n_active) return;
    const int lay = d.active[a];
    const int gi = blockIdx.x * blockDim.x + threadIdx.x;
    if (gi >= d.G) return;
    intensity_group(d, lay, gi, d.pop + (size_t)lay * d.N, false); // the iteration: physics.cuh, dev_intensity_single
}

// every layer (no active list), populations from pop + lay * pop_stride (stride 0: the same vector for all layers)
__global__ void __launch_bounds__(128) k_intensity_all(DevModel d, const double* __restrict__ pop, size_t pop_stride)
{
    const int lay = blockIdx.y;
    const int gi = blockIdx.x * blockDim.x + threadIdx.x;
    if (gi >= d.G) return;
    intensity_group(d, lay, gi, pop + (size_t)lay * pop_stride, true); // one-shot evaluation for lim_luminosity_lvg
}

void launch_intensity(const DevModel& d, cudaStream_t s, int nlay_bound)
{
    if (d.G == 0) return;
    dim3 grid((d.G + 127) / 128, nlay_bound);
    g_rt_launches++;
    rt_launch(k_intensity, grid, 128, 0, s, d, nlay_bound);
}

// ------------------------------------------------------------------------------------------------------------
// K6: loss rates of the two levels of listed transitions, lim_luminosity_lvg
// (maser_luminosity.cpp:44-76).  One CTA per (layer, transition); threads stride over the partner level i:
//   radiative: every A_ij != 0 with i outside the transition and no inversion between i and j (tested with the
//   populations of the CURRENT layer) contributes the out-of-j rate A_ji (1 + J) (i < j) or A_ji J (i > j); the rates
//   are the yd / yu of k_intensity_all (J from the populations the host chose); a pair with A != 0 that is not a line
//   of the scheme (hyperfine doublets of the overlap scheme, iteration_lvg.cpp:503-526) has J = 0;
//   collisional: sum over i != j of get_rate_neutrals(j -> i) (coll_rates.cpp:225-240), neutrals only.
// out [T][L][2] = {low_loss_rate, up_loss_rate}.
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_loss_rates(DevModel d, const double* __restrict__ pop, int nb_trans, const int* __restrict__ up,
                                                    const int* __restrict__ low, double* __restrict__ out)
{
    const int lay = blockIdx.x, t = blockIdx.y;
    const int N = d.N, lo = low[t], hi = up[t];
    const double* __restrict__ p = pop + (size_t)lay * N;
    const double* __restrict__ yd = d.yd + (size_t)lay * d.R;
    const double* __restrict__ yu = d.yu + (size_t)lay * d.R;
    double acc[2] = {0., 0.};
    for (int i = threadIdx.x; i < N; i += blockDim.x) {
#pragma unroll
        for (int k = 0; k < 2; k++) {
            const int j = k == 0 ? lo : hi;
            if (i == j) continue;
            const double Aij = d.einst[(size_t)i * N + j], Aji = d.einst[(size_t)j * N + i];
            if (Aij != 0. && i != lo && i != hi) {
                // the reference compares the two ROUNDED products (maser_luminosity.cpp:53,57): no contraction into
                // fma(p_i, A_ij, -(p_j A_ji)) > 0, which decides differently when the pair is within rounding of balance
                const double lhs = __dmul_rn(p[i], Aij), rhs = __dmul_rn(p[j], Aji);
                const bool down = i < j && lhs > rhs;
                const bool upw = i > j && lhs < rhs;
                if (down || upw) {
                    double rate = down ? Aji : 0.; // J = 0 when the pair is not a line of the scheme
                    for (int e = d.row_ptr[j]; e < d.row_ptr[j + 1]; e++)
                        if (d.row_col[e] == i) {
                            const int ln = d.row_line[e];
                            rate = (ln & 1) ? yu[ln >> 1] : yd[ln >> 1]; // the out-of-row rate, as in k_assemble
                            break;
                        }
                    acc[k] += rate;
                }
            }
            double dn, ur;
            if (j > i) { rate_neutrals(d, lay, j, i, dn, ur); acc[k] += dn; }
            else { rate_neutrals(d, lay, i, j, dn, ur); acc[k] += ur; }
        }
    }
    __shared__ double red[2][4];
#pragma unroll
    for (int k = 0; k < 2; k++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], o);
        if ((threadIdx.x & 31) == 0) red[k][threadIdx.x >> 5] = acc[k];
    }
    __syncthreads();
    if (threadIdx.x < 2) {
        const int k = threadIdx.x;
        out[((size_t)t * d.L + lay) * 2 + k] = red[k][0] + red[k][1] + red[k][2] + red[k][3];
    }
}

void launch_loss_rates(const DevModel& d, cudaStream_t s, const double* pop_dev, const double* pop_intensity, size_t intensity_stride,
                       int nb_trans, const int* up_dev, const int* low_dev, double* out_dev)
{
    if (d.G > 0) {
        dim3 gi((d.G + 127) / 128, d.L);
        g_rt_launches++;
        k_intensity_all<<<gi, 128, 0, s>>>(d, pop_intensity, intensity_stride);
    }
    dim3 grid(d.L, nb_trans);
    g_rt_launches++;
    k_loss_rates<<<grid, 128, 0, s>>>(d, pop_dev, nb_trans, up_dev, low_dev, out_dev);
}

// ------------------------------------------------------------------------------------------------------------
// K1b: M = C + R, row 0 <- 1, df = e0 - M * pop, eq_error = max |df|.
// One warp per matrix row: streams the C row (coalesced 16-byte loads), patches the few radiative entries of
// the row from its sparse list, writes the M row (coalesced), and reduces the row's dot product with pop.
// HBM traffic per layer: 8 Np^2 read + 8 Np^2 write (+ O(R)) -- the algorithmic minimum for a materialised M.
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void atomic_max_nonneg(unsigned long long* addr, double v)
{
    // |df| >= 0: IEEE-754 order of non-negative doubles == order of their bit patterns
    atomicMax(addr, (unsigned long long)__double_as_longlong(v));
}

#define ASM_WARPS 8
__global__ void __launch_bounds__(ASM_WARPS * 32) k_assemble(DevModel d, int nlay_bound, double* __restrict__ df_out)
{
    rt_pdl_sync();
    const int a = blockIdx.y;
    if (a >= nlay_bound || a >= *d.n_active) return;
    const int lay = d.active[a];
    const int Np = d.Np, N = d.N;
    extern __shared__ double s_pop[]; // [Np]
    const double* __restrict__ pop = d.pop + (size_t)lay * N;
    for (int j = threadIdx.x; j < Np; j += blockDim.x) s_pop[j] = j < N ? pop[j] : 0.;
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row = blockIdx.x * ASM_WARPS + warp;
    if (row >= Np) return;
    const double* __restrict__ Crow = d.Cmat + ((size_t)lay * Np + row) * Np;
    double* __restrict__ Mrow = d.Mmat + ((size_t)lay * Np + row) * d.ld;
    // b = e0 (iteration_lvg.cpp:97-98) rides as column Np; columns Np+1 .. ld-1 are zero padding
    if (lane < d.ld - Np) Mrow[Np + lane] = (lane == 0 && row == 0) ? 1. : 0.;
    double dot = 0.;
    if (row == 0) {
        // normalisation row, iteration_lvg.cpp:149-151; padded columns stay 0
        for (int j = lane; j < Np; j += 32) {
            const double v = j < N ? 1. : 0.;
            Mrow[j] = v;
            dot += v * s_pop[j];
        }
    }
    else {
        for (int j = 2 * lane; j < Np; j += 64) {
            const double2 c = *reinterpret_cast<const double2*>(Crow + j);
            *reinterpret_cast<double2*>(Mrow + j) = c;
            dot += c.x * s_pop[j] + c.y * s_pop[j + 1];
        }
        if (row < N) {
            __syncwarp();
            const double* __restrict__ yd = d.yd + (size_t)lay * d.R;
            const double* __restrict__ yu = d.yu + (size_t)lay * d.R;
            double diag = 0.;
            for (int e = d.row_ptr[row] + lane; e < d.row_ptr[row + 1]; e += 32) {
                const int col = d.row_col[e];
                const int ln = d.row_line[e];
                const int line = ln >> 1;
                // kind 0: row is the upper level u, col = l: M[u][l] += yu, M[u][u] -= yd
                // kind 1: row is the lower level l, col = u: M[l][u] += yd, M[l][l] -= yu
                const double off = (ln & 1) ? yd[line] : yu[line];
                const double dg = (ln & 1) ? yu[line] : yd[line];
                Mrow[col] = Crow[col] + off; // each (row, col) pair belongs to exactly one line
                dot += off * s_pop[col];
                diag += dg;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) diag += __shfl_xor_sync(0xffffffffu, diag, o);
            if (lane == 0) {
                Mrow[row] = Crow[row] - diag;
                dot -= diag * s_pop[row];
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
    if (lane == 0 && row < N) {
        const double df = (row == 0 ? 1. : 0.) - dot; // iteration_lvg.cpp:153-160
        if (df_out) df_out[(size_t)lay * N + row] = df;
        atomic_max_nonneg(d.eq_bits + lay, fabs(df));
    }
}

void launch_assemble(const DevModel& d, cudaStream_t s, int nlay_bound, double* df_out)
{
    dim3 grid((d.Np + ASM_WARPS - 1) / ASM_WARPS, nlay_bound);
    g_rt_launches++;
    rt_launch(k_assemble, grid, ASM_WARPS * 32, d.Np * sizeof(double), s, d, nlay_bound, df_out);
}
