// fused_kernels.cu -- the whole iteration of one layer in ONE persistent CTA, for molecules whose rate matrix fits in
// shared memory (OH hyperfine N = 56, NH3, H2CO: register-resident LU; up to N ~ 158, e.g. H2O at the N = 150 the reference
// driver uses: the UNPADDED N x (N + 1) matrix fills the CTA's shared memory and the molecule constants stay in L2).
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

#define FUSED_DYN_SMEM_MAX (210 * 1024) // + up to 13.3 KB static (publication ring) stays under the 227 KB per-CTA limit
#define FUSED_THREADS CTL_THREADS // the controller device code is written for CTL_THREADS cooperating threads

// shared-memory carve-up (doubles): M [Np][ldm] with ldm = Np + 1 (column Np = rhs; odd stride: conflict-free column
// walks), pop [Np], yd [R], yu [R]
__host__ __device__ inline bool fused_c_in_smem(int Np) { return Np <= 96; } // 2 x 73 KB at Np = 96
// constants of the molecule staged once per CTA (read every iteration by the intensity and assembly phases): line records,
// groups, and the per-row sparse lists; doubles first, then the int arrays
__host__ __device__ inline size_t fused_const_doubles(const DevModel& d)
{
    const size_t R = d.R > 0 ? d.R : 1, G = d.G > 0 ? d.G : 1;
    const size_t ints = 3 * G + 2 * R + (size_t)d.N + 1 + 4 * R + 2 * G /* tasks */;
    size_t axes = (size_t)d.lvg.nb_d + d.lvg.nb_g;
    if (d.scheme == RT_SCHEME_LINE_OVERLAP)
        for (int t = 0; t < 2; t++) axes += (size_t)d.ov[t].nb_d + d.ov[t].nb_dx + d.ov[t].nb_gr + d.ov[t].nb_g;
    const size_t halves = d.scheme == RT_SCHEME_LINE_OVERLAP ? 6 * G : 0;
    return 3 * R + R * (d.ncomp > 0 ? d.ncomp : 1) + G + axes + halves + (ints + 1) / 2 + 2;
}
// Dimension of the matrix the CTA factors: N, NOT padded to a multiple of 32 (N = 150: 181 KB instead of 206 KB, the
// difference between fitting and not fitting; OH N = 56: 56 pivot steps instead of 64), rounded up to an even number so that
// the row stride N + 1 stays odd (conflict-free column walks); an odd N gets one identity row / column, which is what Cmat
// holds there anyway.
// (written as N + (N & 1): with (N + 1) & ~1 the compiler proves n even and the register LU of NQ >= 4, which sits at the
// 255-register limit, comes out with 0.3 - 1.9 KB of spills)
__host__ __device__ inline int fused_dim(const DevModel& d) { return d.N + (d.N & 1); }
// cs: molecule constants staged in shared memory (small molecules); otherwise read through L1/L2 every iteration
__host__ __device__ inline size_t fused_smem_doubles(const DevModel& d, bool cs)
{
    const size_t Nm = fused_dim(d), R = d.R > 0 ? d.R : 1;
    return Nm * (Nm + 1) + Nm + 2 * R + 64 + (fused_c_in_smem(d.Np) ? Nm * (Nm + 1) : 0) + (cs ? fused_const_doubles(d) : 0);
}
static bool fused_fits(const DevModel& d, bool cs);
__device__ __forceinline__ void fused_cp_async8(void* smem, const void* gmem)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}

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

// back substitution by one warp, NQ position slots per lane (n <= 32 NQ: lanes beyond n idle along), right-looking, no
// barriers.  Rows are scaled by 1 / U_rr as they are read (off the dependency chain), so that a step of the chain is one
// shuffle (the finished x_i) and one fused multiply-add:  c_r = b_r / U_rr,  c_r -= (U_ri / U_rr) x_i,  x_r = c_r.
// (ncu, OH: with the loads left inside the unrolled chain every step waited for its own shared-memory load: ~120 cycles per
// step, 12 % of the iteration with the other seven warps at the barrier.)
template <int NQ>
__device__ __forceinline__ void fused_backsolve(const double* M, int ldm, int n, const short* rowf, double* gnew, int lane)
{
    double c[NQ], rd[NQ];
    const double* rowp[NQ];
#pragma unroll
    for (int q = 0; q < NQ; q++) {
        const int r = lane + 32 * q;
        const bool ok = r < n;
        rowp[q] = M + (size_t)rowf[ok ? r : 0] * ldm;
        rd[q] = ok ? __drcp_rn(rowp[q][r]) : 0.;
        c[q] = ok ? rowp[q][n] * rd[q] : 0.;
    }
    // eight steps at a time: first the scaled entries of the eight columns (independent loads), then the chain itself
#pragma unroll
    for (int qi = NQ - 1; qi >= 0; qi--) {
#pragma unroll 1
        for (int i0 = 32 * qi + 31; i0 >= 32 * qi; i0 -= 8) {
            if (i0 - 7 >= n) continue; // uniform
            double sc[NQ][8];
            // columns >= n and, in the diagonal slot, the rows at or below the diagonal get a zero: the chain below is
            // branch-free (two shuffles and one multiply-add per step)
#pragma unroll
            for (int t = 0; t < 8; t++) {
                const bool in = i0 - t < n;
                const int i = in ? i0 - t : 0;
#pragma unroll
                for (int q = 0; q < qi; q++) sc[q][t] = in ? __dmul_rn(rowp[q][i], rd[q]) : 0.;
                sc[qi][t] = in && lane < ((i0 - t) & 31) ? __dmul_rn(rowp[qi][i], rd[qi]) : 0.;
            }
#pragma unroll
            for (int t = 0; t < 8; t++) {
                const double xi = __shfl_sync(0xffffffffu, c[qi], (i0 - t) & 31);
#pragma unroll
                for (int q = 0; q <= qi; q++) c[q] -= sc[q][t] * xi;
            }
        }
    }
#pragma unroll
    for (int q = 0; q < NQ; q++)
        if (lane + 32 * q < n) gnew[lane + 32 * q] = c[q];
}

// ------------------------------------------------------------------------------------------------------------------
// Register-resident LU for every dimension n = fused_dim <= 150 (n <= 32 NQ rows, NS = n / 8 + 1 column slots; OH N = 56,
// NH3, H2CO: NQ = 2; o-H2O at the reference driver's N = 150: NQ = 5).  The matrix lives in the registers of the CTA's 8
// warps for the whole factorisation:
//   warp w owns the columns c == w (mod 8)  (slot j <-> column w + 8 j; column n is the right-hand side),
//   lane l owns the rows  l, l + 32, ...     (NQ rows per thread; rows >= n are marked as used and never loaded or stored).
// Step k:  every warp gets the pivot row's values of ITS columns with one shuffle per slot (the pivot row is a lane, not
// memory), the multipliers l_i = a[i][k] / pivot come from a 2 x 32 NQ double-buffered shared array written by the warp that
// owns column k -- that warp ran one column AHEAD during step k-1 (updated column k first, REDUX arg-max on packed keys,
// drcp, multipliers), so the CTA needs ONE barrier per column and no shared-memory traffic for the matrix itself.
// Rows never move: a used-row mask gives the candidates, s_row[k] = pivot row of step k the row order of U at the end.
// Column slots are processed in blocks of 8 columns (one per warp); a finished block is stored to shared memory (U for the
// back substitution) and the slots shift down, so the step body is the same code for every k of a block; a block runs
// min(8, n - 8 KB) steps.
// ------------------------------------------------------------------------------------------------------------------
template <int NQ> __device__ __forceinline__ double sel_row(const double (&v)[NQ], int h)
{
    double r = v[0];
#pragma unroll
    for (int i = 1; i < NQ; i++) r = h == i ? v[i] : r;
    return r;
}
// executed by ONE warp: pivot of column k1 among the rows not yet used as pivots (`done` bit i = row lane + 32 i is used);
// publishes the multipliers of the column and the pivot row.
// Fast path: one REDUX over the top 32 bits of |a| (sign stripped); when exactly one row holds that maximum it is the
// pivot.  Otherwise (two rows agree in 20 mantissa bits, or the column is zero) packed 64-bit keys decide: larger |a| in the
// top 56 bits, exact ties -> smallest row index (the scalar loop takes the smallest position: same magnitude, the choice
// only reorders rounding).
template <int NQ>
__device__ __forceinline__ void reg_pivot(const double (&v)[NQ], unsigned done, int k1, int lane, double* sl, int* spv, short* s_row)
{
    unsigned hmax = 0u;
    int bi = 0;
    bool dup = false;
    double vbest = 0.; // the lane's candidate travels with its key: the winner's value and slot are two independent shuffles
#pragma unroll
    for (int i = 0; i < NQ; i++) {
        const unsigned h = (done >> i) & 1u ? 0u : ((unsigned)__double2hiint(v[i]) & 0x7fffffffu);
        dup = (h == hmax && h != 0u) || (dup && h < hmax);
        if (h > hmax) { hmax = h; bi = i; vbest = v[i]; }
    }
    const unsigned mhi = __reduce_max_sync(0xffffffffu, hmax);
    const unsigned vote = __ballot_sync(0xffffffffu, hmax == mhi);
    const unsigned vdup = __ballot_sync(0xffffffffu, dup && hmax == mhi);
    int wl, wi;
    double piv;
    if (mhi != 0u && __popc(vote) == 1 && vdup == 0u) {
        wl = __ffs(vote) - 1;
        wi = __shfl_sync(0xffffffffu, bi, wl);
        piv = __shfl_sync(0xffffffffu, vbest, wl);
    }
    else {
        unsigned long long mykey[NQ], key = 0ull;
#pragma unroll
        for (int i = 0; i < NQ; i++) {
            mykey[i] = (done >> i) & 1u ? 0ull : pivot_key(v[i], lane + 32 * i);
            key = mykey[i] > key ? mykey[i] : key;
        }
        const unsigned hi = (unsigned)(key >> 32);
        const unsigned khi = __reduce_max_sync(0xffffffffu, hi);
        const unsigned lo = hi == khi ? (unsigned)key : 0u;
        const unsigned klo = __reduce_max_sync(0xffffffffu, lo);
        const unsigned long long win = ((unsigned long long)khi << 32) | klo;
        int myi = -1;
#pragma unroll
        for (int i = 0; i < NQ; i++)
            if (mykey[i] == win) myi = i;
        wl = __ffs(__ballot_sync(0xffffffffu, myi >= 0)) - 1;
        wi = __shfl_sync(0xffffffffu, myi, wl);
        piv = __shfl_sync(0xffffffffu, sel_row<NQ>(v, wi), wl);
    }
    const int p1 = wl + 32 * wi;
    const double rp = piv != 0. ? __drcp_rn(piv) : 0.; // 0: zero pivot, the column is left as is
#pragma unroll
    for (int i = 0; i < NQ; i++) {
        const int row = lane + 32 * i;
        sl[row] = ((done >> i) & 1u) || row == p1 ? 0. : v[i] * rp;
    }
    if (lane == 0) { spv[0] = p1; s_row[k1] = (short)p1; }
}

// One LU step on the column slots [J0, NACT) of every warp
template <int NQ, int NS, int J0, int NACT>
__device__ __forceinline__ void reg_update(double (&a)[NQ][NS], const double (&l)[NQ], int ph, int pl)
{
#pragma unroll
    for (int j = J0; j < NACT; j++) {
        double col[NQ];
#pragma unroll
        for (int i = 0; i < NQ; i++) col[i] = a[i][j];
        const double u = __shfl_sync(0xffffffffu, sel_row<NQ>(col, ph), pl);
#pragma unroll
        for (int i = 0; i < NQ; i++) a[i][j] -= l[i] * u;
    }
}

// Steps 8 KB .. 8 KB + 7 (fewer in the last block), one CTA barrier per step; the publications (multipliers + pivot row)
// are double-buffered by column parity.  Measured alternatives, all slower or equal on B200 (OH N = 56, per iteration): warps decoupled through
// per-step mbarriers so that the rank-1 updates trail behind the pivot chain: 30 us vs 31 us; spinning on a shared-memory
// counter instead of mbarriers: 155 us; speculative reciprocals of every lane's candidate: slower; columns BLOCKED by warps
// (warp w owns 8 consecutive columns and factors its sub-panel alone, warp-synchronously; the others apply the 8 pivots after
// ONE barrier: 8 barriers per factorisation instead of 64, bitwise the same factors): 71.1 vs 72.0 ms per OH solve, i.e.
// equal -- ncu then shows the owner warp's factorisation alone covering 74 % of the wall time at ~180 instructions and
// ~700 cycles per column.  The step is bound by the instruction count of the warp that searches the pivot (~3-4 cycles per
// dependent instruction of a single warp), not by the barrier.  The same two questions asked again at N = 150 (NQ = 5, 19
// slots, where the rank-1 updates are 10 x the work): publications through an 8-deep ring guarded by full / empty mbarriers
// instead of the CTA barrier: 18.4 vs 18.6 ms per solve; a warp-uniform switch on the pivot row's slot instead of the 5-way
// select per column slot (a third of the step's instructions): 18.5 vs 18.6 ms -- equal, not kept.
template <int NQ, int NS, int KB>
__device__ __forceinline__ void reg_lu_block(double (&a)[NQ][NS], unsigned& done, double* M, int ldm, int n, short* s_row, double* s_l,
                                              int (*s_pv)[4], int w, int lane)
{
    constexpr int Np = 32 * NQ, NACT = NS - KB;
    if (8 * KB > n) return; // uniform: column n (the right-hand side) was the last one
    const int nsteps = n - 8 * KB < 8 ? n - 8 * KB : 8;
    for (int kk = 0; kk < nsteps; kk++) {
        const int k = 8 * KB + kk, par = k & 1;
        const int p = s_pv[par][0];
        double l[NQ];
#pragma unroll
        for (int i = 0; i < NQ; i++) {
            l[i] = s_l[par * Np + lane + 32 * i];
            done |= (unsigned)(lane + 32 * i == p) << i;
        }
        const int ph = p >> 5, pl = p & 31;
        {
            double col[NQ];
#pragma unroll
            for (int i = 0; i < NQ; i++) col[i] = a[i][0];
            const double u = __shfl_sync(0xffffffffu, sel_row<NQ>(col, ph), pl);
            if (w > kk) {
#pragma unroll
                for (int i = 0; i < NQ; i++) a[i][0] -= l[i] * u;
            }
        }
        if (NACT > 1) reg_update<NQ, NS, 1, (NACT > 1 ? 2 : 1)>(a, l, ph, pl);
        if (k + 1 < n && w == ((kk + 1) & 7)) {
            double v[NQ];
#pragma unroll
            for (int i = 0; i < NQ; i++) v[i] = (kk == 7 && NACT > 1) ? a[i][NACT > 1 ? 1 : 0] : a[i][0];
            reg_pivot<NQ>(v, done, k + 1, lane, s_l + (par ^ 1) * Np, s_pv[par ^ 1], s_row);
        }
        if (NACT > 2) reg_update<NQ, NS, 2, NACT>(a, l, ph, pl);
        __syncthreads();
    }
    if (w + 8 * KB <= n) {
#pragma unroll
        for (int i = 0; i < NQ; i++)
            if (lane + 32 * i < n) M[(size_t)(lane + 32 * i) * ldm + w + 8 * KB] = a[i][0];
    }
#pragma unroll
    for (int i = 0; i < NQ; i++) {
#pragma unroll
        for (int j = 0; j + 1 < NACT; j++) a[i][j] = a[i][j + 1];
    }
}
template <int NQ, int NS, int KB> struct RegLuBlocks {
    static __device__ __forceinline__ void run(double (&a)[NQ][NS], unsigned& done, double* M, int ldm, int n, short* s_row, double* s_l,
                                               int (*s_pv)[4], int w, int lane)
    {
        reg_lu_block<NQ, NS, KB>(a, done, M, ldm, n, s_row, s_l, s_pv, w, lane);
        RegLuBlocks<NQ, NS, KB + 1>::run(a, done, M, ldm, n, s_row, s_l, s_pv, w, lane);
    }
};
template <int NQ, int NS> struct RegLuBlocks<NQ, NS, NS> {
    static __device__ __forceinline__ void run(double (&)[NQ][NS], unsigned&, double*, int, int, short*, double*, int (*)[4], int, int) {}
};
// n <= 8 NS - 1 and n <= 32 NQ (checked by the launcher).  All FUSED_THREADS threads call it; ends with a barrier.
template <int NQ, int NS>
__device__ __forceinline__ void fused_lu_regs(double* M, int ldm, int n, short* s_row, double* s_l, int (*s_pv)[4], int w, int lane)
{
    double a[NQ][NS];
    unsigned done = 0u;
#pragma unroll
    for (int i = 0; i < NQ; i++) {
        const int row = lane + 32 * i;
        const double* Mr = M + (size_t)row * ldm;
        done |= (unsigned)(row >= n) << i;
#pragma unroll
        for (int j = 0; j < NS; j++) a[i][j] = (row < n && w + 8 * j <= n) ? Mr[w + 8 * j] : 0.;
    }
    if (w == 0) {
        double v[NQ];
#pragma unroll
        for (int i = 0; i < NQ; i++) v[i] = a[i][0];
        reg_pivot<NQ>(v, done, 0, lane, s_l, s_pv[0], s_row);
    }
    __syncthreads();
    RegLuBlocks<NQ, NS, 0>::run(a, done, M, ldm, n, s_row, s_l, s_pv, w, lane);
    __syncthreads();
}

// column slots of the register LU with NQ row slots: n / 8 + 1 for the largest n of the class (32, 62, 96, 128, 150)
__host__ __device__ constexpr int fused_lu_slots(int nq) { return nq == 1 ? 5 : nq == 2 ? 8 : nq == 3 ? 13 : nq == 4 ? 17 : 19; }
__host__ __device__ inline int fused_lu_variant(int n) { return n <= 32 ? 1 : n <= 62 ? 2 : n <= 96 ? 3 : n <= 128 ? 4 : n <= 150 ? 5 : 0; }

// LUV: 1..5 = register-resident LU with NQ = LUV row slots (n <= 32, 62, 96, 128, 150), 0 = LU in shared memory (n <= 160).
// One kernel per variant so that the register allocation of the small, common sizes (two CTAs per SM) is not set by the
// large ones.
// CS: molecule constants staged in shared memory (false: the matrix needs the room; they are read through L1/L2).
template <int LUV, bool CS = true>
__global__ void __launch_bounds__(FUSED_THREADS, (LUV == 1 || LUV == 2) ? 2 : 1)
k_fused_solve(DevModel d, int nlay, int accel, int max_iter, double min_err)
{
    const int a = blockIdx.x;
    if (a >= nlay || a >= *d.n_active) return;
    const int lay = d.active[a];
    // Np: dimension of the matrix in shared memory (fused_dim); NpG: padded dimension of the global arrays (Cmat, rhs)
    const int N = d.N, Np = fused_dim(d), NpG = d.Np, ldm = Np + 1, R = d.R;
    extern __shared__ __align__(16) double sm[];
    double* M = sm;                      // [Np][ldm]
    double* s_pop = M + (size_t)Np * ldm; // [Np]
    double* s_yd = s_pop + Np;           // [R]
    double* s_yu = s_yd + (R > 0 ? R : 1);
    __shared__ short s_row[160];  // position -> physical row of the LU
    __shared__ FusedPiv s_piv[2]; // pivot of the column, double-buffered by column parity
    __shared__ double s_l[2 * (LUV > 0 ? 32 * LUV : 1)]; // register-resident LU: multipliers of the column, double-buffered
    __shared__ int s_pv[2][4];                           //   pivot row of the column
    __shared__ double s_eq[FUSED_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31;
    // The warp index through a broadcast shuffle: the compiler then knows it is warp-uniform, keeps it in a uniform register
    // and drops the divergence guards (BRA.DIV + reconvergence) it otherwise puts in front of every warp collective of the
    // pivot search, which sits under `if (warp == owner)`: 38 -> 3 guards in the N = 56 kernel, OH 65.6 -> 59.7 ms per solve.
    // (The same line in the kernels of the blocked LU made them slower -- 4.45 vs 4.33 ms per CH3OH iteration -- and is not
    // used there.)
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    // the collisional matrix of the layer is constant over the iteration: staged in shared memory when it fits
    const double* C = d.Cmat + (size_t)lay * NpG * NpG;
    int ldc = NpG;
    double* arena = s_yu + (R > 0 ? R : 1);
    if (fused_c_in_smem(Np)) {
        double* s_C = arena;
        ldc = Np + 1; // odd stride: the assembly reads 8 rows per warp
        for (int e = tid; e < Np * Np; e += FUSED_THREADS) s_C[(e / Np) * ldc + e % Np] = C[(size_t)(e / Np) * NpG + e % Np];
        C = s_C;
        arena += (size_t)Np * ldc;
    }
    // molecule constants: staged in shared memory (CS, fused_const_doubles) or read in place through L1/L2
    const int Rn = R > 0 ? R : 1, Gn = d.G > 0 ? d.G : 1, nc = d.ncomp > 0 ? d.ncomp : 1, nnz = d.row_ptr[N];
    const double *c_Aul = d.tr_Aul, *c_Alu = d.tr_Alu, *c_E = d.tr_E, *c_sig = d.tr_sigma, *c_dxn = d.grp_dxn;
    const int *c_grp = d.grp, *c_tru = d.tr_u, *c_trl = d.tr_l, *c_rptr = d.row_ptr, *c_rcol = d.row_col, *c_rline = d.row_line;
    const int* c_task = nullptr; // [2G] group * 2 + half; without staging (LVG scheme only) task t is line t
    double* c_half = nullptr;    // [2G][3] halves of the pair evaluations (overlap scheme)
    DevLvg lvg = d.lvg;
    DevOverlap ov0 = d.ov[0], ov1 = d.ov[1];
    __shared__ int s_ntask;
    if constexpr (CS) {
        double* w_Aul = arena; double* w_Alu = w_Aul + Rn; double* w_E = w_Alu + Rn; double* w_sig = w_E + Rn;
        double* w_dxn = w_sig + (size_t)Rn * nc;
        // axes of the escape-probability tables (binary searches every iteration)
        double* a = w_dxn + Gn;
        for (int e = tid; e < lvg.nb_d; e += FUSED_THREADS) a[e] = d.lvg.delta[e];
        lvg.delta = a; a += lvg.nb_d;
        for (int e = tid; e < lvg.nb_g; e += FUSED_THREADS) a[e] = d.lvg.gamma[e];
        lvg.gamma = a; a += lvg.nb_g;
        if (d.scheme == RT_SCHEME_LINE_OVERLAP) {
#pragma unroll
            for (int t = 0; t < 2; t++) {
                DevOverlap& o = t ? ov1 : ov0;
                const DevOverlap& g = d.ov[t];
                for (int e = tid; e < o.nb_d; e += FUSED_THREADS) a[e] = g.log_delta[e];
                o.log_delta = a; a += o.nb_d;
                for (int e = tid; e < o.nb_dx; e += FUSED_THREADS) a[e] = g.dx[e];
                o.dx = a; a += o.nb_dx;
                for (int e = tid; e < o.nb_gr; e += FUSED_THREADS) a[e] = g.gratio[e];
                o.gratio = a; a += o.nb_gr;
                for (int e = tid; e < o.nb_g; e += FUSED_THREADS) a[e] = g.gamma[e];
                o.gamma = a; a += o.nb_g;
            }
        }
        c_half = a;
        int* w_grp = (int*)(c_half + (d.scheme == RT_SCHEME_LINE_OVERLAP ? 6 * Gn : 0)); int* w_tru = w_grp + 3 * Gn; int* w_trl = w_tru + Rn;
        int* w_rptr = w_trl + Rn; int* w_rcol = w_rptr + N + 1; int* w_rline = w_rcol + 2 * Rn; int* w_task = w_rline + 2 * Rn;
        for (int e = tid; e < R; e += FUSED_THREADS) {
            w_Aul[e] = d.tr_Aul[e]; w_Alu[e] = d.tr_Alu[e]; w_E[e] = d.tr_E[e]; w_tru[e] = d.tr_u[e]; w_trl[e] = d.tr_l[e];
            for (int k = 0; k < d.ncomp; k++) w_sig[(size_t)e * d.ncomp + k] = d.tr_sigma[(size_t)e * d.ncomp + k];
        }
        for (int e = tid; e < d.G; e += FUSED_THREADS) {
            w_dxn[e] = d.grp_dxn[e];
            w_grp[3 * e] = d.grp[3 * e]; w_grp[3 * e + 1] = d.grp[3 * e + 1]; w_grp[3 * e + 2] = d.grp[3 * e + 2];
        }
        if (d.scheme != RT_SCHEME_LINE_OVERLAP) { // one task per line
            for (int g = tid; g < d.G; g += FUSED_THREADS) w_task[g] = 2 * g;
            if (tid == 0) s_ntask = d.G;
        }
        else {
            __syncthreads();
            if (tid == 0) { // one task per single line, two per overlapping pair
                int nt = 0;
                for (int g = 0; g < d.G; g++) {
                    w_task[nt++] = 2 * g;
                    if (w_grp[3 * g] == 2) w_task[nt++] = 2 * g + 1;
                }
                s_ntask = nt;
            }
        }
        for (int e = tid; e <= N; e += FUSED_THREADS) w_rptr[e] = d.row_ptr[e];
        for (int e = tid; e < nnz; e += FUSED_THREADS) { w_rcol[e] = d.row_col[e]; w_rline[e] = d.row_line[e]; }
        c_Aul = w_Aul; c_Alu = w_Alu; c_E = w_E; c_sig = w_sig; c_dxn = w_dxn;
        c_grp = w_grp; c_tru = w_tru; c_trl = w_trl; c_rptr = w_rptr; c_rcol = w_rcol; c_rline = w_rline; c_task = w_task;
    }
    else if (tid == 0) s_ntask = d.G; // LVG scheme only (fused_supported): one task per line
    const bool same_axes = d.ov_same_axes != 0;
    double* gpop = d.pop + (size_t)lay * N;
    double* gnew = d.rhs + (size_t)lay * NpG;
    const LayerCtx ctx = dev_layer_ctx(d, lay);

    for (;;) {
        // ---- populations of this step (the controller left them in global memory) ----------------------------------
        for (int j = tid; j < Np; j += FUSED_THREADS) s_pop[j] = j < N ? gpop[j] : 0.;
        // N > 96 (C is read from global memory / L2): the copy M = C starts here as 8-byte asynchronous copies -- M is free
        // since the back substitution -- and lands while the intensities are computed; the assembly then patches in place
        constexpr bool C_ASYNC = !(LUV >= 1 && LUV <= 3);
        if (C_ASYNC && !fused_c_in_smem(Np)) {
            for (int r = warp; r < Np; r += FUSED_THREADS / 32)
                for (int c = lane; c < Np; c += 32) fused_cp_async8(M + (size_t)r * ldm + c, C + (size_t)r * ldc + c);
            asm volatile("cp.async.commit_group;" ::: "memory");
        }
        __syncthreads();
        // ---- K1a: mean intensities -> radiative rates (iteration_lvg.cpp:163-185, :428-501) ------------------------
        const int ntask = s_ntask;
        for (int t = tid; t < ntask; t += FUSED_THREADS) {
            const int task = CS ? c_task[t] : 2 * t;
            const int gi = task >> 1, half = task & 1;
            const int nb = c_grp[3 * gi], la = c_grp[3 * gi + 1];
            const int u1 = c_tru[la], l1 = c_trl[la];
            const double A1ul = c_Aul[la], A1lu = c_Alu[la];
            if (nb == 1) {
                const double J = dev_intensity_single(d, lvg, ctx, c_E[la], A1ul, A1lu, s_pop[u1], s_pop[l1], c_sig + (size_t)la * d.ncomp);
                s_yd[la] = A1ul * (1. + J);
                s_yu[la] = A1lu * J;
            }
            else {
                const int lb = c_grp[3 * gi + 2];
                const int u2 = c_tru[lb], l2 = c_trl[lb];
                const PairCommon pc = dev_pair_common(d, ctx, c_E[la], c_dxn[gi], A1ul, A1lu, c_Aul[lb], c_Alu[lb], s_pop[u1], s_pop[l1],
                                                      s_pop[u2], s_pop[l2], c_sig + (size_t)la * d.ncomp);
                double e[3];
                dev_pair_half(d, lvg, ov0, ov1, same_axes, pc, half, e);
                double* h = c_half + 3 * (size_t)(2 * gi + half);
                h[0] = e[0]; h[1] = e[1]; h[2] = e[2];
            }
        }
        if (d.scheme == RT_SCHEME_LINE_OVERLAP) {
            __syncthreads();
            for (int gi = tid; gi < d.G; gi += FUSED_THREADS) {
                if (c_grp[3 * gi] != 2) continue;
                const int la = c_grp[3 * gi + 1], lb = c_grp[3 * gi + 2];
                const int u1 = c_tru[la], l1 = c_trl[la], u2 = c_tru[lb], l2 = c_trl[lb];
                const double A1ul = c_Aul[la], A1lu = c_Alu[la], A2ul = c_Aul[lb], A2lu = c_Alu[lb];
                const PairCommon pc = dev_pair_common(d, ctx, c_E[la], c_dxn[gi], A1ul, A1lu, A2ul, A2lu, s_pop[u1], s_pop[l1],
                                                      s_pop[u2], s_pop[l2], c_sig + (size_t)la * d.ncomp);
                const double* h = c_half + 6 * (size_t)gi;
                const double a[3] = {h[0], h[1], h[2]}, b[3] = {h[3], h[4], h[5]};
                double J1, J2;
                dev_pair_combine(d, pc, a, b, J1, J2);
                s_yd[la] = A1ul * (1. + J1); s_yu[la] = A1lu * J1;
                s_yd[lb] = A2ul * (1. + J2); s_yu[lb] = A2lu * J2;
            }
        }
        const bool m_is_c = C_ASYNC && !fused_c_in_smem(Np); // M holds C already (asynchronous copy above)
        if (m_is_c) asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();
        // ---- K1b: M = C + R, row 0 = 1, rhs = e0, df = e0 - M pop (iteration_lvg.cpp:112-161); one warp per row ------
        // four lanes per row: 64 rows in flight, two shuffle steps per reduction (a whole warp per row -- coalesced loads of
        // C where it is read from L2, N > 96 -- was measured at N = 150: 168 vs 151 us per iteration, slower)
        double eqmax = 0.;
        {
            const int q = tid & 3;
            const unsigned gmask = 0xFu << (lane & ~3);
            for (int row = tid >> 2; row < Np; row += FUSED_THREADS / 4) {
                double* Mrow = M + (size_t)row * ldm;
                const double* Crow = C + (size_t)row * ldc;
                double dot = 0.;
                if (row == 0) {
                    for (int j = q; j < Np; j += 4) { const double v = j < N ? 1. : 0.; Mrow[j] = v; dot += v * s_pop[j]; }
                }
                else {
                    if (m_is_c) {
                        for (int j = q; j < Np; j += 4) dot += Mrow[j] * s_pop[j];
                    }
                    else {
                        for (int j = q; j < Np; j += 4) { const double c = Crow[j]; Mrow[j] = c; dot += c * s_pop[j]; }
                    }
                    if (row < N) {
                        __syncwarp(gmask); // the row's copy is complete before its radiative entries are patched
                        double diag = 0.;
                        for (int e = c_rptr[row] + q; e < c_rptr[row + 1]; e += 4) {
                            const int col = c_rcol[e], ln = c_rline[e], line = ln >> 1;
                            const double off = (ln & 1) ? s_yd[line] : s_yu[line];
                            const double dg = (ln & 1) ? s_yu[line] : s_yd[line];
                            Mrow[col] = (m_is_c ? Mrow[col] : Crow[col]) + off;
                            dot += off * s_pop[col];
                            diag += dg;
                        }
                        diag += __shfl_xor_sync(gmask, diag, 1);
                        diag += __shfl_xor_sync(gmask, diag, 2);
                        if (q == 0) { Mrow[row] = (m_is_c ? Mrow[row] : Crow[row]) - diag; dot -= diag * s_pop[row]; }
                    }
                }
                dot += __shfl_xor_sync(gmask, dot, 1);
                dot += __shfl_xor_sync(gmask, dot, 2);
                if (q == 0) {
                    Mrow[Np] = row == 0 ? 1. : 0.;
                    if (row < N) eqmax = fmax(eqmax, fabs((row == 0 ? 1. : 0.) - dot));
                }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) eqmax = fmax(eqmax, __shfl_xor_sync(0xffffffffu, eqmax, o));
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
        if constexpr (LUV > 0) fused_lu_regs<LUV, fused_lu_slots(LUV)>(M, ldm, Np, s_row, s_l, s_pv, warp, lane);
        else {
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
            if constexpr (LUV > 0 && LUV <= 3) fused_backsolve<LUV>(M, ldm, Np, s_row, gnew, lane);
            else if (Np <= 128) fused_backsolve<4>(M, ldm, Np, s_row, gnew, lane);
            else fused_backsolve<5>(M, ldm, Np, s_row, gnew, lane);
        }
        __syncthreads();
        // ---- K3: controller step (iteration_control.h:84-137), shared with the multi-kernel path --------------------
        if (ctl_step_body(d, lay, accel, max_iter, min_err)) break;
        __syncthreads();
    }
}

// RT_B200_FUSED_SHARED_LU=1: the shared-memory LU for every Np > 96 (the round-2 path, kept for A/B measurements and as the
// path of 151 < n <= 158)
static int g_fused_shared_lu = 0;
static bool fused_fits(const DevModel& d, bool cs) { return fused_smem_doubles(d, cs) * sizeof(double) <= FUSED_DYN_SMEM_MAX; }

void launch_fused_solve(const DevModel& d, cudaStream_t s, int nlay, int accel, int max_iter, double min_err)
{
    g_rt_launches++;
    const bool cs = fused_fits(d, true);
    const size_t smem = fused_smem_doubles(d, cs) * sizeof(double);
    const int n = fused_dim(d);
    int v = fused_lu_variant(n);
    if (v >= 4 && g_fused_shared_lu) v = 0;
    switch (v) {
    case 1: k_fused_solve<1><<<nlay, FUSED_THREADS, smem, s>>>(d, nlay, accel, max_iter, min_err); break;
    case 2: k_fused_solve<2><<<nlay, FUSED_THREADS, smem, s>>>(d, nlay, accel, max_iter, min_err); break;
    case 3: k_fused_solve<3><<<nlay, FUSED_THREADS, smem, s>>>(d, nlay, accel, max_iter, min_err); break;
    case 4:
        if (cs) k_fused_solve<4, true><<<nlay, FUSED_THREADS, smem, s>>>(d, nlay, accel, max_iter, min_err);
        else k_fused_solve<4, false><<<nlay, FUSED_THREADS, smem, s>>>(d, nlay, accel, max_iter, min_err);
        break;
    case 5:
        if (cs) k_fused_solve<5, true><<<nlay, FUSED_THREADS, smem, s>>>(d, nlay, accel, max_iter, min_err);
        else k_fused_solve<5, false><<<nlay, FUSED_THREADS, smem, s>>>(d, nlay, accel, max_iter, min_err);
        break;
    default:
        if (cs) k_fused_solve<0, true><<<nlay, FUSED_THREADS, smem, s>>>(d, nlay, accel, max_iter, min_err);
        else k_fused_solve<0, false><<<nlay, FUSED_THREADS, smem, s>>>(d, nlay, accel, max_iter, min_err);
    }
}

// One CTA per layer with the whole matrix in shared memory: every size whose UNPADDED matrix fits (N up to ~158 for a
// molecule with ~700 lines).  Without room for the staged constants only the LVG scheme is supported (the overlap scheme's
// pair halves live in that staging area; its molecules are small anyway).
bool fused_supported(const DevModel& d)
{
    if (fused_dim(d) > 160) return false;
    if (d.Np <= 96) return fused_fits(d, true);
    return fused_fits(d, true) || (d.scheme != RT_SCHEME_LINE_OVERLAP && fused_fits(d, false));
}

template <int LUV, bool CS> static cudaError_t fused_set_smem()
{
    return cudaFuncSetAttribute(k_fused_solve<LUV, CS>, cudaFuncAttributeMaxDynamicSharedMemorySize, FUSED_DYN_SMEM_MAX);
}

cudaError_t fused_kernels_init()
{
    cudaError_t e = fused_set_smem<0, true>();
    if (e == cudaSuccess) e = fused_set_smem<0, false>();
    if (e == cudaSuccess) e = fused_set_smem<1, true>();
    if (e == cudaSuccess) e = fused_set_smem<2, true>();
    if (e == cudaSuccess) e = fused_set_smem<3, true>();
    if (e == cudaSuccess) e = fused_set_smem<4, true>();
    if (e == cudaSuccess) e = fused_set_smem<4, false>();
    if (e == cudaSuccess) e = fused_set_smem<5, true>();
    if (e == cudaSuccess) e = fused_set_smem<5, false>();
    const char* v = getenv("RT_B200_FUSED_SHARED_LU");
    g_fused_shared_lu = v && atoi(v) != 0;
    return e;
}
