// lu_kernels.cu -- K2: batched FP64 dense solve  M x = e0, one matrix per active layer.
//
// Replaces lu_matrix_solve (missing from the reference; contract restated in oracle/shim/linear_algebra.h;
// call site iteration_lvg.cpp:97-101): LU with ROW PARTIAL PIVOTING (largest |a_ik| of the column, first
// occurrence wins), multipliers l = a_ik * (1/pivot), right-hand side eliminated on the fly, back substitution.
//
// Layout: Mmat [L][Np][ld] row-major, ld = Np + 32; column Np is the right-hand side (so the forward substitution
// is just one more column of every update), columns Np+1.. are zero padding.
//
// Two-level blocked right-looking algorithm, inner panel 32 columns, outer block 64 columns:
//   for every outer block at k0:
//     k_panel (k0)        factor the (Np-k0) x 32 panel, REGISTER-resident (each thread owns up to 3 rows x 32
//                         columns), pivot search by warp shuffles; records the composed row permutation
//     k_trsm  (k0, mode 1) columns right of the panel: apply the permutation, U1 = L11^-1 A
//     k_gemm  (K = 32)    update only the next 32 columns (the second panel of the block)
//     k_panel (k0+32)     second panel; also permutes the L block of the first panel (LAPACK-style) so that
//     k_trsm  (k0+32, mode 2) rows k0+32..k0+63 of the columns right of the block: permutation, -= L21 U1, L22^-1
//     k_gemm  (K = 64)    trailing update A22 -= [L1 L2] [U1; U2] on FP64 tensor cores (DMMA mma.sync m8n8k4):
//                         the O(N^3) part; 8 flop per byte of trailing matrix moved instead of 4 at K = 32
//   k_backsolve           U x = y, blocked from the bottom.
// Work per layer: (2/3) Np^3 + O(Np^2) flop; HBM traffic ~ sum over outer blocks of 16 m^2 bytes.
#include "rt_internal.h"
#include <cfloat>
#include <cstdlib>
#include <algorithm>

#define PNL_MAXROWS 768 // largest panel height = largest Np
#ifndef PNL_BIG_R
#define PNL_BIG_R 3     // rows per thread x threads of the kernel variant for 512 < m <= 768
#define PNL_BIG_NT 256
#endif
#ifndef PNL_JB
#define PNL_JB 8   // columns per unrolled block of the panel's column loop
#endif

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem)
{
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void cp_async8(void* smem, const void* gmem)
{
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(s), "l"(gmem));
}

// ------------------------------------------------------------------------------------------------------------
// Panel factorisation, register-resident and blocked inside the panel.  Local row r = t + NT i (i < R) lives in the
// registers x[i][0..31] of thread t.  Per column j: (1) arg-max |x[.][j]| over rows >= j (ties -> smallest row, like the
// scalar loop), two-stage reduction with ONE barrier; (2) the owner of the pivot row publishes it through shared
// memory -- second barrier; (3) rank-1 update in registers.  Shared buffers are double-buffered by the parity of j.
// An unblocked version (every column step updates all 32 columns of every row) was bound by the instructions a warp
// issues between two barriers (ncu: ~500 cycles of fixed latency + ~3.4 cycles per instruction with 2 warps per
// scheduler; 400 instructions per column at 3 rows per thread, of which 82 DFMA on the 32-wide row and 48 LDS of the
// pivot row; 62 us per m = 768 panel against 45 us now).  Measured alternatives, not kept: one barrier per column (each
// warp's local winner publishes its candidate row; equal), 384 threads x 2 rows for m > 512 (equal), rows loaded straight
// into registers instead of through the staging tile (6 % slower), 512 threads x 1 row for 256 < m <= 512 (4 % slower).
// Round 2, measured and not kept: the SECOND panel of an outer block fused with what precedes it on the critical chain (the
// near-column TRSM and the 32-deep next-panel update): rows loaded through the first panel's permutation, U12 by 32 threads,
// A22 -= L21 U12 inside the staging tile on DMMA (bitwise the same factors).  It removes two kernels per outer block, but
// serialises on ONE SM per layer work that those kernels spread over the whole GPU: 8 layers 1.13 -> 1.18 ms per
// iteration, 200 layers 4.32 -> 4.40 ms (panel class 1.19 -> 1.52 ms, update class 2.53 -> 2.29 ms).  The same update by
// plain FMAs on the register-resident rows was worse still (every thread reads all 8 KB of U12: +40 us per panel).
// For 256 < m <= 384 the kernel runs as 192 threads x 2 rows: 168 registers keep TWO CTAs per SM, so 200 layers are one
// wave instead of two (1.21 -> 1.16 ms per N = 768 slab-iteration, 0.46 -> 0.41 ms at N = 411).
// Here the 32 columns are factored as four SUB-PANELS of 8 columns: inside a
// sub-panel a column step touches only the 8 sub-panel columns of every row (<= 7 DFMA per row, 4 LDS, the pivot
// row published as 8 doubles); the 24 (16, 8) columns to the right are updated once per sub-panel by a dense rank-8
// update from registers -- no barrier, no pivot search in between:
//     U12 = L11^-1 A12      24 threads, one column each (the 8 pivot rows publish their trailing parts)
//     A22 -= L21 U12        every thread on its own rows, multipliers L21 = its x[i][0..7]
// For every matrix entry this is the same sequence of fused multiply-adds, in the same order, as the eager rank-1
// updates: the factors are bitwise identical to the unblocked kernel's.
// The window of live columns is shifted left by 8 registers after every sub-panel (register indices must be
// static; one code body serves the four sub-panels), finished entries go to the staging tile.
// ------------------------------------------------------------------------------------------------------------
// Rows that are pivots already (and padding slots) are updated too: their registers are dead, nobody reads them again.
template <int R, int NC>
__device__ __forceinline__ void panel_rank8(double (&x)[R][32], const double* __restrict__ u12 /*[8][24]*/)
{
#pragma unroll
    for (int c = 0; c < NC; c += 2) {
#pragma unroll
        for (int q = 0; q < 8; q++) {
            const double2 u = *reinterpret_cast<const double2*>(u12 + q * 24 + c);
#pragma unroll
            for (int i = 0; i < R; i++) {
                x[i][8 + c] -= x[i][q] * u.x;
                x[i][9 + c] -= x[i][q] * u.y;
            }
        }
    }
}

template <int R, int NT>
__global__ void __launch_bounds__(NT, (R * NT <= 384 ? 2 : 1)) k_panel(DevModel d, int k0, int nleft, int a0, int a1)
{
    rt_pdl_sync();
    const int a = a0 + blockIdx.x;
    if (a >= a1 || a >= *d.n_active) return;
    const int lay = d.active[a];
    const int Np = d.Np, ld = d.ld, m = Np - k0;
    double* __restrict__ A = d.Mmat + (size_t)lay * Np * ld;
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;

    extern __shared__ double s_tile[]; // [m][33]: coalesced staging of the panel on the way in and out
    __shared__ unsigned long long s_key[2][NT / 32];
    __shared__ __align__(16) double s_prow[2][8];  // sub-panel columns of the pivot row
    __shared__ __align__(16) double s_l11[8][8];   // sub-panel columns of the 8 pivot rows (their multipliers: unit-lower L11)
    __shared__ __align__(16) double s_a12[8][24];  // trailing columns of the 8 pivot rows, overwritten by U12
    __shared__ int s_ownrow[8];                    // local row (t + NT i) that holds pivot q of the sub-panel
    __shared__ int s_pos[64], s_cur[64], s_naff;   // composed permutation (see k_trsm)
    __shared__ short s_slot[NT * R]; // local row -> slot in s_pos, -1 = untouched

    // panel -> staging tile with 8-byte cp.async (the tile's row stride of 33 doubles is not 16-byte aligned): every
    // thread has all of its m / 8 copies in flight at once
    for (int r = warp; r < m; r += NT / 32) cp_async8(s_tile + r * 33 + lane, A + (size_t)(k0 + r) * ld + k0 + lane);
    asm volatile("cp.async.commit_group;" ::: "memory");
    for (int r = t; r < m; r += NT) s_slot[r] = (short)(r < 32 ? r : -1);
    if (t < 32) { s_pos[t] = t; s_cur[t] = t; }
    if (t == 0) s_naff = 32;
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();

    // Rows never move between threads: vp[i] is the CURRENT position of the row held in x[i] (the position the
    // scalar algorithm's interchanges would have put it at), -1 for padding slots.  x[i][c] holds column jb + c.
    double x[R][32];
    int vp[R];
#pragma unroll
    for (int i = 0; i < R; i++) {
        const int r = t + NT * i;
        vp[i] = r < m ? r : -1;
#pragma unroll
        for (int c = 0; c < 32; c++) x[i][c] = r < m ? s_tile[r * 33 + c] : 0.;
    }

#pragma unroll 1
    for (int jb = 0; jb < 32; jb += 8) {
#pragma unroll
        for (int jj = 0; jj < 8; jj++) {
            const int j = jb + jj;
            const int buf = jj & 1;
            // ---- (1) pivot search over the rows at positions >= j.  64-bit keys: |a| with its low 10 bits replaced by
            // 1023 - (position - j): the larger value wins, values equal in their top 54 bits (relative difference
            // < 2e-13, immaterial for the pivot choice) count as ties -> smallest position, like the scalar loop.
            // Warp arg-max = two REDUX instructions instead of five shuffle rounds.
            unsigned long long key = 0ull;
#pragma unroll
            for (int i = 0; i < R; i++) {
                if (vp[i] >= j) {
                    const unsigned long long kk = ((unsigned long long)__double_as_longlong(fabs(x[i][jj])) & ~0x3FFull)
                                                  | (unsigned long long)(1023 - (vp[i] - j));
                    key = kk > key ? kk : key;
                }
            }
            {
                const unsigned hi = (unsigned)(key >> 32);
                const unsigned mhi = __reduce_max_sync(0xffffffffu, hi);
                const unsigned lo = hi == mhi ? (unsigned)key : 0u;
                const unsigned mlo = __reduce_max_sync(0xffffffffu, lo);
                if (lane == 0) s_key[buf][warp] = ((unsigned long long)mhi << 32) | mlo;
            }
            __syncthreads();
            {   // max over the warps' keys: one load per lane + two REDUX instead of NT/32 loads and a compare chain
                const unsigned long long kk = lane < NT / 32 ? s_key[buf][lane] : 0ull;
                const unsigned hi = (unsigned)(kk >> 32);
                const unsigned mhi = __reduce_max_sync(0xffffffffu, hi);
                const unsigned lo = hi == mhi ? (unsigned)kk : 0u;
                const unsigned mlo = __reduce_max_sync(0xffffffffu, lo);
                key = ((unsigned long long)mhi << 32) | mlo;
            }
            const int p = j + 1023 - (int)(key & 0x3FFull);
            // ---- (2) interchange of positions j <-> p: the pivot row publishes its sub-panel columns, nothing else moves
#pragma unroll
            for (int i = 0; i < R; i++) {
                if (vp[i] == p) {
                    double2* pr = reinterpret_cast<double2*>(s_prow[buf]);
#pragma unroll
                    for (int c = 0; c < 4; c++) pr[c] = make_double2(x[i][2 * c], x[i][2 * c + 1]);
                    s_ownrow[jj] = t + NT * i;
                    vp[i] = j;
                }
                else if (vp[i] == j) vp[i] = p;
            }
            if (t == 0 && p != j) {
                // bookkeeping of the composed permutation: s_cur[slot] = slot of the ORIGINAL row now at s_pos[slot]
                int sp = s_slot[p];
                if (sp < 0) { sp = s_naff++; s_pos[sp] = p; s_cur[sp] = sp; s_slot[p] = (short)sp; }
                const int tmp = s_cur[j]; s_cur[j] = s_cur[sp]; s_cur[sp] = tmp;
            }
            __syncthreads();
            // ---- (3) rank-1 update of the sub-panel columns of the rows below j; the multiplier stays in the register
            const double piv = s_prow[buf][jj];
            const double rp = piv != 0. ? __drcp_rn(piv) : 0.; // == 1. / piv (correctly rounded); zero pivot: column left alone
#pragma unroll
            for (int i = 0; i < R; i++) {
                if (vp[i] > j) {
                    const double l = x[i][jj] * rp;
                    x[i][jj] = l;
#pragma unroll
                    for (int c = jj + 1; c < 8; c++) x[i][c] -= l * s_prow[buf][c];
                }
            }
        }
        // ---- (4) the columns right of the sub-panel: U12 = L11^-1 A12, A22 -= L21 U12
        const int ntr = 24 - jb; // live columns right of the sub-panel (window columns 8 .. 8 + ntr)
        if (ntr > 0) {
#pragma unroll
            for (int i = 0; i < R; i++) {
                const int q = vp[i] - jb;
                if (q >= 0 && q < 8) {
                    double2* l11 = reinterpret_cast<double2*>(s_l11[q]);
#pragma unroll
                    for (int c = 0; c < 4; c++) l11[c] = make_double2(x[i][2 * c], x[i][2 * c + 1]);
                    double2* a12 = reinterpret_cast<double2*>(s_a12[q]);
#pragma unroll
                    for (int c = 0; c < 12; c++) a12[c] = make_double2(x[i][8 + 2 * c], x[i][9 + 2 * c]);
                }
            }
            __syncthreads();
            if (t < ntr) {
                double u[8];
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    double v = s_a12[q][t];
#pragma unroll
                    for (int r = 0; r < q; r++) v -= s_l11[q][r] * u[r];
                    u[q] = v;
                }
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    s_a12[q][t] = u[q];
                    s_tile[s_ownrow[q] * 33 + jb + 8 + t] = u[q]; // final U entries of the pivot rows
                }
            }
            __syncthreads();
            if (jb == 0) panel_rank8<R, 24>(x, &s_a12[0][0]);
            else if (jb == 8) panel_rank8<R, 16>(x, &s_a12[0][0]);
            else panel_rank8<R, 8>(x, &s_a12[0][0]);
        }
        // ---- (5) finished sub-panel columns (multipliers; U entries of the rows that became pivots here) -> staging tile;
        // rows that became pivots in an EARLIER sub-panel are complete there already (their registers are dead)
#pragma unroll
        for (int i = 0; i < R; i++) {
            if (vp[i] >= jb) {
                double* dst = s_tile + (t + NT * i) * 33 + jb;
#pragma unroll
                for (int c = 0; c < 8; c++) dst[c] = x[i][c];
            }
#pragma unroll
            for (int c = 0; c < 24; c++) x[i][c] = x[i][c + 8];
#pragma unroll
            for (int c = 24; c < 32; c++) x[i][c] = 0.;
        }
    }

    // ---- write the factored panel back: row at final position vp <- staging-tile row of its owner (coalesced)
    __shared__ short s_owner[NT * R];
#pragma unroll
    for (int i = 0; i < R; i++)
        if (vp[i] >= 0) s_owner[vp[i]] = (short)(t + NT * i);
    __syncthreads();
    for (int r = warp; r < m; r += NT / 32) A[(size_t)(k0 + r) * ld + k0 + lane] = s_tile[s_owner[r] * 33 + lane];
    const int naff = s_naff;
    int* aff = d.lu_aff + ((size_t)lay * (Np / RT_NB) + k0 / RT_NB) * RT_AFF_STRIDE;
    if (t == 0) aff[0] = naff;
    if (t < 64) { aff[4 + t] = t < naff ? s_pos[t] : 0; aff[68 + t] = t < naff ? s_cur[t] : 0; }
    // ---- LAPACK-style interchange of the `nleft` columns left of the panel (the L block of the first panel of
    // this outer block): gather every affected row, barrier, scatter
    if (nleft > 0) {
        // one warp per (affected row, 32-column chunk): nleft <= 96 (fourth panel of a super-block), naff <= 64
        constexpr int NQ = (64 * 3 + NT / 32 - 1) / (NT / 32);
        double v[NQ];
        const int w = nleft >> 5, nrc = naff * w;
#pragma unroll
        for (int q = 0; q < NQ; q++) {
            const int rc = warp + (NT / 32) * q;
            if ((NT / 32) * q >= nrc) break; // uniform
            if (rc < nrc) {
                const int s = w == 1 ? rc : (w == 2 ? rc >> 1 : (rc * 171) >> 9), ch = rc - s * w;
                v[q] = A[(size_t)(k0 + s_pos[s_cur[s]]) * ld + k0 - nleft + 32 * ch + lane];
            }
        }
        __syncthreads();
#pragma unroll
        for (int q = 0; q < NQ; q++) {
            const int rc = warp + (NT / 32) * q;
            if ((NT / 32) * q >= nrc) break; // uniform
            if (rc < nrc) {
                const int s = w == 1 ? rc : (w == 2 ? rc >> 1 : (rc * 171) >> 9), ch = rc - s * w;
                if (s_cur[s] != s) A[(size_t)(k0 + s_pos[s]) * ld + k0 - nleft + 32 * ch + lane] = v[q];
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------------------
// Row permutation + (deferred left update) + triangular solve for the columns right of a panel.  One thread per
// column, 128 columns per CTA.  The panel kernel left the COMPOSED permutation of its 32 interchanges: naff affected
// local rows pos[0..naff) (pos[s] = s for s < 32, then the pivot rows taken from below) and cur[s] = slot of the
// original row that ends at pos[s].  Rows displaced to below the block always come from the top 32.
//   x = rows k0..k0+31 of the column (after the permutation)
//   x -= Lleft u      Lleft = A[k0..k0+31][k0-nl..k0-1] (the L entries of the pivot rows left of this panel inside the
//                     current outer block, already permuted by k_panel), u = rows k0-nl..k0-1 of the column (final U);
//                     nl = nleft_lo for the columns left of c_split, nleft_hi from c_split on (columns inside a
//                     128-wide super-block have received part of these updates from a restricted trailing update)
//   x <- L11^-1 x
// The left update runs on the FP64 tensor cores: per warp a (32 x 32) x (32 x 32) product per 32 rows of u (staged
// by the warp's own lanes for its own 32 columns: only warp-level synchronisation), 128 DMMA + 64 fragment loads
// instead of 1024 DFMA + 544 shared-memory loads per thread (a scalar sweep is bound by the shared-memory pipe).
// The product goes back through the warp's slice of s_u to the one-column-per-thread layout of the substitution.
// ------------------------------------------------------------------------------------------------------------
#define TRSM_THREADS 128
#ifndef TRSM_LEFT_MINB
#define TRSM_LEFT_MINB 3
#endif
#define TRSM_LDT 34 // even: rows of the transposed triangles stay 16-byte aligned for double2 loads
#define TRSM_LDU (TRSM_THREADS + 4) // row stride of s_u: = 4 (mod 16) doubles -> conflict-free DMMA fragment loads
static size_t trsm_smem_bytes(int nleft) { return nleft > 0 ? (size_t)(32 * TRSM_LDU + nleft * TRSM_LDT) * sizeof(double) : 0; }

template <bool LEFT> // LEFT = false: no pending left update anywhere (first panel of a block): 128 registers, 4 CTAs/SM
__global__ void __launch_bounds__(TRSM_THREADS, LEFT ? TRSM_LEFT_MINB : 4) k_trsm(DevModel d, int k0, int c0, int c_end, int nleft_lo, int nleft_hi, int c_split, int a0, int a1)
{
    rt_pdl_sync();
    const int a = a0 + blockIdx.y;
    if (a >= a1 || a >= *d.n_active) return;
    const int lay = d.active[a];
    const int Np = d.Np, ld = d.ld;
    __shared__ __align__(16) double Lt[32 * TRSM_LDT];  // Lt[q][r] = L[r][q] of the panel's unit-lower 32x32 block
    __shared__ int s_pos[64], s_cur[64];
    extern __shared__ __align__(16) double s_dyn[];
    double* s_u = s_dyn;                  // [32][TRSM_LDU] 32 rows of u for this column tile, then the product
    double* Llt = s_dyn + 32 * TRSM_LDU;  // [nleft_hi][TRSM_LDT]: Llt[q][r] = A[k0+r][k0-nleft_hi+q]
    double* __restrict__ A = d.Mmat + (size_t)lay * Np * ld;
    const int* aff = d.lu_aff + ((size_t)lay * (Np / RT_NB) + k0 / RT_NB) * RT_AFF_STRIDE;
    const int naff = aff[0];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int r = warp; r < 32; r += TRSM_THREADS / 32) {
        Lt[lane * TRSM_LDT + r] = A[(size_t)(k0 + r) * ld + k0 + lane];
        if (LEFT) for (int q = lane; q < nleft_hi; q += 32) Llt[q * TRSM_LDT + r] = A[(size_t)(k0 + r) * ld + k0 - nleft_hi + q];
    }
    if (tid < 64) { s_pos[tid] = aff[4 + tid]; s_cur[tid] = aff[68 + tid]; }
    __syncthreads();
    const int c = c0 + blockIdx.x * TRSM_THREADS + tid;
    if (c >= c_end) return; // c0 and c_end are multiples of 32: whole warps leave
    const int nl = !LEFT ? 0 : ((c - lane) < c_split ? nleft_lo : nleft_hi); // warp-uniform
    const int fr = lane >> 2, fk = lane & 3;
    if (LEFT && nl > 0) {
        // the pivot rows are needed only after the left update: lane r pulls row r of the warp's 256-byte column slice
        // towards the SM meanwhile
        {
            const double* q = A + (size_t)(k0 + s_pos[s_cur[lane]]) * ld + (c - lane);
            asm volatile("prefetch.global.L1 [%0];" ::"l"(q));
            asm volatile("prefetch.global.L1 [%0];" ::"l"(q + 16));
        }
        // P = Lleft u accumulated over chunks of 32 rows of u; the next chunk's loads are in flight during the DMMAs
        double acc[2][4][2][2];
#pragma unroll
        for (int h = 0; h < 2; h++)
#pragma unroll
            for (int ti = 0; ti < 4; ti++)
#pragma unroll
                for (int tj = 0; tj < 2; tj++) acc[h][ti][tj][0] = acc[h][ti][tj][1] = 0.;
        double un[32];
#pragma unroll
        for (int q = 0; q < 32; q++) un[q] = A[(size_t)(k0 - nl + q) * ld + c];
        double* Bw = s_u + fk * TRSM_LDU + warp * 32;
        for (int ch = 0; ch < nl; ch += 32) {
#pragma unroll
            for (int q = 0; q < 32; q++) s_u[q * TRSM_LDU + tid] = un[q];
            __syncwarp();
            if (ch + 32 < nl) {
#pragma unroll
                for (int q = 0; q < 32; q++) un[q] = A[(size_t)(k0 - nl + ch + 32 + q) * ld + c];
            }
            const double* Aw = Llt + (nleft_hi - nl + ch + fk) * TRSM_LDT + fr;
#pragma unroll
            for (int k4 = 0; k4 < 8; k4++) {
                double af[4], bf[4];
#pragma unroll
                for (int ti = 0; ti < 4; ti++) af[ti] = Aw[4 * k4 * TRSM_LDT + ti * 8];
#pragma unroll
                for (int tj = 0; tj < 4; tj++) bf[tj] = Bw[4 * k4 * TRSM_LDU + tj * 8 + fr];
#pragma unroll
                for (int ti = 0; ti < 4; ti++)
#pragma unroll
                    for (int tj = 0; tj < 4; tj++) dmma884(acc[tj >> 1][ti][tj & 1][0], acc[tj >> 1][ti][tj & 1][1], af[ti], bf[tj]);
            }
            __syncwarp(); // every lane has read its fragments: the next chunk (or the product) may overwrite the slice
        }
#pragma unroll
        for (int h = 0; h < 2; h++)
#pragma unroll
            for (int ti = 0; ti < 4; ti++)
#pragma unroll
                for (int tj = 0; tj < 2; tj++)
                    *reinterpret_cast<double2*>(s_u + (ti * 8 + fr) * TRSM_LDU + warp * 32 + h * 16 + tj * 8 + 2 * fk) = make_double2(acc[h][ti][tj][0], acc[h][ti][tj][1]);
        __syncwarp();
    }
    // rows k0..k0+31 after the permutation: gathered straight from global memory (the same row for every thread of
    // the CTA -> coalesced); all gathers precede the stores of this column, columns are independent
    double x[32];
#pragma unroll
    for (int r = 0; r < 32; r++) x[r] = A[(size_t)(k0 + s_pos[s_cur[r]]) * ld + c];
    // rows displaced below the block receive ORIGINAL top rows (still untouched in memory: they are rewritten last)
    for (int s = 32; s < naff; s++)
        if (s_cur[s] != s) A[(size_t)(k0 + s_pos[s]) * ld + c] = A[(size_t)(k0 + s_pos[s_cur[s]]) * ld + c];
    if (LEFT && nl > 0) {
#pragma unroll
        for (int r = 0; r < 32; r++) x[r] -= s_u[r * TRSM_LDU + tid];
    }
    // forward substitution, right-looking: after x[q] is final, every x[r], r > q, is updated independently
#pragma unroll
    for (int q = 0; q < 31; q++) {
        const double2* lq = reinterpret_cast<const double2*>(Lt + q * TRSM_LDT);
#pragma unroll
        for (int r2 = (q + 1) / 2; r2 < 16; r2++) {
            const double2 l = lq[r2];
            if (2 * r2 > q) x[2 * r2] -= l.x * x[q];
            x[2 * r2 + 1] -= l.y * x[q];
        }
    }
#pragma unroll
    for (int r = 0; r < 32; r++) A[(size_t)(k0 + r) * ld + c] = x[r];
}

// ------------------------------------------------------------------------------------------------------------
// Trailing update on the FP64 tensor cores:  C[i0.., c0..c1) -= A[i0.., kk0..kk0+K) * A[kk0..kk0+K, c0..c1),
// K = 32 or 64.  CTA tile 64 x 64, 4 warps (2 x 2), warp tile 32 x 32 = 4 x 4 DMMA m8n8k4 tiles (SASS DMMA.8x8x4; the
// m16n8k8 / m16n8k16 f64 shapes of mma.sync compile to the same DMMA.8x8x4 sequence on sm_100a -- checked with cuobjdump --
// so they save no issue slots), accumulators initialised with C; D = C - A*B is accumulated by feeding the NEGATED A
// fragment.  A and B tiles are staged in shared memory with 16-byte cp.async, a 64-deep update as two 32-deep stages
// through the same buffer: four CTAs per SM overlap one CTA's loads and stores with the others' DMMA stream.
// Shared-memory strides are = 4 (mod 16) doubles so that every fragment load is bank-conflict free:
//   A fragment: lane holds As[row = lane/4][k = lane%4]   (row-major tile, stride K+4)
//   B fragment: lane holds Bs[k = lane%4][col = lane/4]   (row-major tile, stride 68)
// ------------------------------------------------------------------------------------------------------------
#define GM_COLS 64
#define GM_LDB 68

// NST stages of K columns each (kk0, kk0 + K, ...): the accumulators stay in registers across the stages, so C is read and
// written once per NST * K eliminated columns while one stage's A / B tiles fit four CTAs per SM.
// COLS = 32: the tile of the NEXT-PANEL update (32 columns wide): two warps, 64 x 32, so that no half of a 64-wide tile
// multiplies zeros; half the threads, registers and shared memory per CTA.  With the driver's per-kernel carve-out this
// third kernel made the grouped iteration SLOWER (4.40 -> 4.72 ms, as did a 128 x 32 tile); with the common carve-out set in
// lu_kernels_init it is faster (4.36 -> 4.32 ms).
template <int K, int ROWS, int NST, int MINB = (ROWS == 64 ? 3 : 2), int COLS = GM_COLS>
__global__ void __launch_bounds__(ROWS * COLS / 32, MINB) k_gemm(DevModel d, int i0, int kk0, int c0, int c1, int a0, int a1)
{
    constexpr int NT = ROWS * COLS / 32;            // one warp per 32 x 32 block of the tile
    constexpr int LDB = COLS == 64 ? GM_LDB : 36;   // row stride of the B tile, = 4 (mod 16) doubles
    rt_pdl_sync();
    const int a = a0 + blockIdx.z;
    if (a >= a1 || a >= *d.n_active) return;
    const int lay = d.active[a];
    const int Np = d.Np, ld = d.ld;
    constexpr int LDA = K + 4;
    extern __shared__ __align__(16) double sm[];
    double* As = sm;                 // [ROWS][LDA]
    double* Bs = sm + ROWS * LDA; // [K][68]
    double* __restrict__ A = d.Mmat + (size_t)lay * Np * ld;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int r0 = i0 + blockIdx.y * ROWS, cc0 = c0 + blockIdx.x * COLS;
    // ---- stage A (rows r0.., K columns at kk) and B (K rows at kk, 64 columns at cc0): 16-byte cp.async
    constexpr int ACH = K / 2; // 16-byte chunks per A row
    auto stage = [&](int kk) {
        for (int e = tid; e < ROWS * ACH; e += NT) {
            const int r = e / ACH, ch = e - r * ACH;
            double* dst = As + r * LDA + 2 * ch;
            if (r0 + r < Np) cp_async16(dst, A + (size_t)(r0 + r) * ld + kk + 2 * ch);
            else { dst[0] = 0.; dst[1] = 0.; }
        }
        for (int e = tid; e < K * (COLS / 2); e += NT) {
            const int q = e / (COLS / 2), ch = e % (COLS / 2);
            double* dst = Bs + q * LDB + 2 * ch;
            if (cc0 + 2 * ch < c1) cp_async16(dst, A + (size_t)(kk + q) * ld + cc0 + 2 * ch);
            else { dst[0] = 0.; dst[1] = 0.; }
        }
        asm volatile("cp.async.commit_group;");
    };
    stage(kk0);
    // ---- accumulators <- C
    const int wr = warp / (COLS / 32), wc = warp % (COLS / 32);
    const int rw = r0 + wr * 32, cw = cc0 + wc * 32;
    const int fr = lane >> 2, fc = (lane & 3) * 2;
    double acc[4][4][2];
#pragma unroll
    for (int ti = 0; ti < 4; ti++)
#pragma unroll
        for (int tj = 0; tj < 4; tj++) {
            const int r = rw + ti * 8 + fr, c = cw + tj * 8 + fc;
            if (r < Np && c < c1) {
                const double2 v = *reinterpret_cast<const double2*>(A + (size_t)r * ld + c);
                acc[ti][tj][0] = v.x; acc[ti][tj][1] = v.y;
            }
            else acc[ti][tj][0] = acc[ti][tj][1] = 0.;
        }
    const double* Aw = As + (wr * 32 + fr) * LDA + (lane & 3);
    const double* Bw = Bs + (lane & 3) * LDB + wc * 32 + fr;
#pragma unroll 1
    for (int st = 0; st < NST; st++) {
        if (st > 0) {
            __syncthreads(); // every warp is done with the previous stage's tiles
            stage(kk0 + st * K);
        }
        asm volatile("cp.async.wait_group 0;");
        __syncthreads();
#pragma unroll 4
        for (int k4 = 0; k4 < K / 4; k4++) {
            double af[4], bf[4];
#pragma unroll
            for (int ti = 0; ti < 4; ti++) af[ti] = -Aw[ti * 8 * LDA + k4 * 4];
#pragma unroll
            for (int tj = 0; tj < 4; tj++) bf[tj] = Bw[k4 * 4 * LDB + tj * 8];
#pragma unroll
            for (int ti = 0; ti < 4; ti++)
#pragma unroll
                for (int tj = 0; tj < 4; tj++) dmma884(acc[ti][tj][0], acc[ti][tj][1], af[ti], bf[tj]);
        }
    }
#pragma unroll
    for (int ti = 0; ti < 4; ti++)
#pragma unroll
        for (int tj = 0; tj < 4; tj++) {
            const int r = rw + ti * 8 + fr, c = cw + tj * 8 + fc;
            if (r < Np && c < c1) *reinterpret_cast<double2*>(A + (size_t)r * ld + c) = make_double2(acc[ti][tj][0], acc[ti][tj][1]);
        }
}

// ------------------------------------------------------------------------------------------------------------
// U x = y (y = column Np after the elimination), blocked from the bottom; one CTA per layer.  x -> d.rhs [L][Np].
// ------------------------------------------------------------------------------------------------------------
#define BS_THREADS 512
__global__ void __launch_bounds__(BS_THREADS) k_backsolve(DevModel d, int a0, int a1)
{
    rt_pdl_sync();
    const int a = a0 + blockIdx.x;
    if (a >= a1 || a >= *d.n_active) return;
    const int lay = d.active[a];
    const int Np = d.Np, ld = d.ld, nb = Np / RT_NB;
    extern __shared__ double ys[]; // [Np]: y, progressively reduced to y - U[:, done] x[done]
    __shared__ double Us[32][33];
    __shared__ __align__(16) double xb[32];
    const double* __restrict__ A = d.Mmat + (size_t)lay * Np * ld;
    double* __restrict__ xo = d.rhs + (size_t)lay * Np;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int r = tid; r < Np; r += BS_THREADS) ys[r] = A[(size_t)r * ld + Np];
    // The update of block kb reads the column panel U[0 .. r0)[r0 .. r0 + 32) right after the triangular solve: a dependent
    // chain of (global-load latency + solve) per block.  The addresses do not depend on x, so the panel (and the diagonal
    // block) of block kb - 1 is prefetched one block ahead, two 128-byte lines per row.
    auto prefetch_block = [&](int kbp) {
        if (kbp < 0) return;
        const int rp = kbp * RT_NB;
        for (int e = tid; e < 2 * (rp + 32); e += BS_THREADS) {
            const double* q = A + (size_t)(e >> 1) * ld + rp + 16 * (e & 1);
            asm volatile("prefetch.global.L1 [%0];" ::"l"(q));
        }
    };
    prefetch_block(nb - 1);
    for (int kb = nb - 1; kb >= 0; kb--) {
        const int r0 = kb * RT_NB;
        prefetch_block(kb - 1);
        for (int rr = warp; rr < 32; rr += BS_THREADS / 32) Us[rr][lane] = A[(size_t)(r0 + rr) * ld + r0 + lane];
        __syncthreads();
        if (warp == 0) {
            // 32 x 32 upper-triangular solve; the scalar loop's divisions become multiplications by reciprocals that
            // all lanes compute at once (the dependent chain per unknown is shuffle + multiply + fma)
            double s = ys[r0 + lane];
            const double diag = Us[lane][lane], rdiag = __drcp_rn(diag);
            double xv = 0.;
            for (int c = 31; c >= 0; c--) {
                // s / diag to the accuracy of a division, as the reference's `b[i] = s / ri[i]` (reciprocal multiply + one
                // residual correction; the plain product s * (1/diag) carries ~1.5 ulp).  Two more FMAs per unknown on the
                // chain: no measurable cost (0.23 ms per N = 768 slab-iteration either way)
                double q = s * rdiag;
                q = fma(fma(-q, diag, s), rdiag, q);
                const double xc = __shfl_sync(0xffffffffu, q, c);
                if (lane == c) xv = xc;
                if (lane < c) s -= Us[lane][c] * xc;
            }
            xb[lane] = xv;
            xo[r0 + lane] = xv;
        }
        __syncthreads();
        // right-looking update of everything above: ys[r] -= U[r][r0..r0+31] . x_blk.  Four threads per row, 8 columns
        // (64 bytes) each; 128 rows per pass, two passes in flight per thread
        const int q4 = tid & 3, rsub = tid >> 2;
        const double2* xq = reinterpret_cast<const double2*>(xb + 8 * q4);
        const double2 x0 = xq[0], x1 = xq[1], x2 = xq[2], x3 = xq[3];
        for (int rbase = 0; rbase < r0; rbase += 2 * (BS_THREADS / 4)) {
            const int ra = rbase + rsub, rb2 = ra + BS_THREADS / 4;
            double va = 0., vb = 0.;
            double2 a0, a1, a2, a3, b0, b1, b2, b3;
            if (ra < r0) {
                const double2* p = reinterpret_cast<const double2*>(A + (size_t)ra * ld + r0 + 8 * q4);
                a0 = p[0]; a1 = p[1]; a2 = p[2]; a3 = p[3];
            }
            if (rb2 < r0) {
                const double2* p = reinterpret_cast<const double2*>(A + (size_t)rb2 * ld + r0 + 8 * q4);
                b0 = p[0]; b1 = p[1]; b2 = p[2]; b3 = p[3];
            }
            if (ra < r0) va = a0.x * x0.x + a0.y * x0.y + a1.x * x1.x + a1.y * x1.y + a2.x * x2.x + a2.y * x2.y + a3.x * x3.x + a3.y * x3.y;
            if (rb2 < r0) vb = b0.x * x0.x + b0.y * x0.y + b1.x * x1.x + b1.y * x1.y + b2.x * x2.x + b2.y * x2.y + b3.x * x3.x + b3.y * x3.y;
            va += __shfl_xor_sync(0xffffffffu, va, 1); vb += __shfl_xor_sync(0xffffffffu, vb, 1);
            va += __shfl_xor_sync(0xffffffffu, va, 2); vb += __shfl_xor_sync(0xffffffffu, vb, 2);
            if (q4 == 0) {
                if (ra < r0) ys[ra] -= va;
                if (rb2 < r0) ys[rb2] -= vb;
            }
        }
        // the next iteration's barrier after the Us load orders these ys updates before warp 0 reads them
    }
}

// ------------------------------------------------------------------------------------------------------------
// host driver
// ------------------------------------------------------------------------------------------------------------
static int g_panel_192 = 1; // development switch RT_B200_PANEL_192
// Every launch works on the slice [a0, a1) of the compacted active list.
static void launch_panel(const DevModel& d, cudaStream_t s, int k0, int nleft, int a0, int a1)
{
    const int m = d.Np - k0, n = a1 - a0;
    g_rt_launches++;
    const size_t smem = (size_t)m * 33 * sizeof(double);
    if (m <= 256) rt_launch(k_panel<1, 256>, n, 256, smem, s, d, k0, nleft, a0, a1);
    else if (m <= 384 && g_panel_192) rt_launch(k_panel<2, 192>, n, 192, smem, s, d, k0, nleft, a0, a1); // 170 registers: still two CTAs per SM
    else if (m <= 512) rt_launch(k_panel<2, 256>, n, 256, smem, s, d, k0, nleft, a0, a1);
    else rt_launch(k_panel<PNL_BIG_R, PNL_BIG_NT>, n, PNL_BIG_NT, smem, s, d, k0, nleft, a0, a1);
}

// 64-deep trailing update: 64 x 64 tiles, TWO 32-deep stages through one 36 KB buffer, 128 registers -> FOUR CTAs/SM
// (2.55 ms per N = 768 slab-iteration).  Measured and removed (same step): 128 x 64 tiles, one 64-deep stage, 2 CTAs/SM
// 2.97 ms; 64 x 64, one stage, 3 CTAs/SM 2.87 ms; five CTAs/SM at 96 registers (spills) 2.64 ms; 32 x 16 warp tiles, eight
// warps per CTA 2.62 - 2.69 ms; B tile resident over two row tiles (-13 % L2 traffic, 53 KB) 2.88 ms; a persistent
// cp.async double-buffered kernel with C prefetch 3.6 ms; K cut into 32-deep stages double-buffered inside the CTA at
// 3 CTAs/SM 3.14 ms; 128-wide super-blocks (third blocking level: 128-deep update 2.47 ms but +0.27 ms TRSM, +0.04 ms
// panel: 5.42 vs 5.52 ms per full-batch iteration, slower with few active layers); round 2: the C tile staged by cp.async
// (accumulators start at zero, the warp never waits for C: long-scoreboard stalls 1.8 -> 0.8 per issue, but 73 KB -> 3 CTAs/SM
// and the DMMA sub-pipe 74.6 % -> 64.4 % busy, profiles/r02_ncu_full_gemm_cstaged.csv): 2.91 ms -- in every
// case the fourth resident CTA is worth more than what the variant saves; `prefetch.global.L2` of the C tile of the CTA that
// runs 148 / 296 / 592 / 1184 block indices later (so that a tile's C loads hit L2 instead of HBM): 2.58 ms at every
// distance against 2.53 ms without -- the update is not waiting for C.  The NEXT-PANEL update (32 columns) on 128 x 32 tiles
// instead of 64 x 64 tiles of which half multiply zeros: 2.54 vs 2.55 ms for the class alone, but 4.73 vs 4.40 ms per
// grouped full-batch iteration (half as many, twice as long CTAs on the critical chain of the look-ahead schedule): removed.
// Layer groups of one solve (development switch RT_B200_LU_GROUPS).  The layers of a batch are independent, so the
// active list is cut into G slices whose factorisations run on G streams: while one slice sits in a latency-bound
// panel (one register-file-filling CTA per SM, 200 layers = 1.35 waves on 148 SMs) the trailing updates of the other
// slices fill the idle SMs.
static int g_gemm32_narrow = 1; // RT_B200_GEMM32_NARROW=0: the next-panel update on 64 x 64 tiles (half of every tile idle)
static int g_lu_groups = 2;
// Measured and removed (round 2): a STAGGERED start of the groups (group g starts when group g-1 leaves its first panel
// phase, so that one group's panels would run beside the other's trailing updates): 4.34 -> 4.39 ms per full-batch
// iteration with 2 groups, 4.41 / 4.45 ms with 3 / 4 -- the idle tail of the late group costs more than the overlap gains.
static int g_lu_lookahead = 1; // development switch RT_B200_LU_LOOKAHEAD=0: every group on one stream
static int g_lu_group_min = 48; // layers per slice below which a slice no longer fills its share of the machine: fewer groups

// K = 32: next panel's columns; 64: trailing update of a 64-wide block
static void launch_gemm(const DevModel& d, cudaStream_t s, int K, int i0, int kk0, int c0, int c1, int a0, int a1)
{
    const int n = a1 - a0;
    const dim3 g64((c1 - c0 + GM_COLS - 1) / GM_COLS, (d.Np - i0 + 63) / 64, n);
    const size_t smem = (64 * 36 + 32 * GM_LDB) * sizeof(double);
    g_rt_launches++;
    if (K == 32 && c1 - c0 <= 32 && g_gemm32_narrow)
        rt_launch(k_gemm<32, 64, 1, 8, 32>, g64, 64, (64 * 36 + 32 * 36) * sizeof(double), s, d, i0, kk0, c0, c1, a0, a1);
    else if (K == 32) rt_launch(k_gemm<32, 64, 1, 4>, g64, 128, smem, s, d, i0, kk0, c0, c1, a0, a1);
    else rt_launch(k_gemm<32, 64, 2, 4>, g64, 128, smem, s, d, i0, kk0, c0, c1, a0, a1);
}

void launch_lu_solve(rt_model* m, cudaStream_t s0, int bound)
{
    const DevModel& d = m->d;
    const int Np = d.Np, ld = d.ld;
    const int c_panel = rt_prof_class(m, "lu_panel"), c_trsm = rt_prof_class(m, "lu_trsm");
    const int c_gemm = rt_prof_class(m, "lu_gemm"), c_back = rt_prof_class(m, "lu_backsolve");
    // per-class event timing needs the launches of a class back to back on one stream: one group when profiling
    int G = m->prof_timing ? 1 : std::min(g_lu_groups, std::max(1, bound / g_lu_group_min));
    G = std::min(G, RT_LU_MAX_GROUPS);
    if (G > 1 && !m->lu_streams[1]) {
        for (int g = 1; g < RT_LU_MAX_GROUPS; g++) {
            if (cudaStreamCreateWithPriority(&m->lu_streams[g], cudaStreamNonBlocking, m->prio_hi) != cudaSuccess ||
                cudaEventCreateWithFlags(&m->lu_join[g], cudaEventDisableTiming) != cudaSuccess) { G = 1; break; }
        }
        if (G > 1 && cudaEventCreateWithFlags(&m->lu_fork, cudaEventDisableTiming) != cudaSuccess) G = 1;
    }
    cudaStream_t st[RT_LU_MAX_GROUPS];
    int ga0[RT_LU_MAX_GROUPS], ga1[RT_LU_MAX_GROUPS];
    for (int g = 0; g < G; g++) {
        st[g] = g == 0 ? s0 : m->lu_streams[g];
        ga0[g] = (int)((long long)bound * g / G);
        ga1[g] = (int)((long long)bound * (g + 1) / G);
    }
    if (G > 1) {
        cudaEventRecord(m->lu_fork, s0);
        for (int g = 1; g < G; g++) cudaStreamWaitEvent(st[g], m->lu_fork, 0);
    }
    const int layers_all = m->prof_layers;
    auto panel = [&](int g, int k0, int nleft) {
        const double mr = Np - k0;
        rt_prof_begin(m, c_panel, st[g], 32. * 32. * mr, 16. * 32. * mr);
        launch_panel(d, st[g], k0, nleft, ga0[g], ga1[g]);
        rt_prof_end(m, c_panel, st[g]);
    };
    // columns [c0, c1) right of the panel at k0: left update nleft_lo wide left of c_split, nleft_hi wide from c_split on
    auto trsm_on = [&](cudaStream_t ss, int g, int k0, int c0, int c1, int nleft_lo, int nleft_hi, int c_split) {
        const double nc = c1 - c0, ncl = std::max(0, std::min(c_split, c1) - c0), nch = nc - ncl;
        // algorithmic columns only: of the 32 columns of padding right of the matrix one (the right-hand side) is real
        const double pad = c1 > Np + 1 ? c1 - (Np + 1) : 0., nca = nc - pad, ncha = nch - pad;
        rt_prof_begin(m, c_trsm, ss, 32. * 32. * nca + 2. * 32. * (nleft_lo * ncl + nleft_hi * ncha),
                      16. * 32. * nca + 8. * (nleft_lo * ncl + nleft_hi * ncha));
        g_rt_launches++;
        const dim3 grid((c1 - c0 + TRSM_THREADS - 1) / TRSM_THREADS, ga1[g] - ga0[g]);
        if (nleft_hi > 0) rt_launch(k_trsm<true>, grid, TRSM_THREADS, trsm_smem_bytes(nleft_hi), ss, d, k0, c0, c1, nleft_lo, nleft_hi, c_split, ga0[g], ga1[g]);
        else rt_launch(k_trsm<false>, grid, TRSM_THREADS, 0, ss, d, k0, c0, c1, 0, 0, c_split, ga0[g], ga1[g]);
        rt_prof_end(m, c_trsm, ss);
    };
    auto trsm = [&](int g, int k0, int nleft_lo, int nleft_hi, int c_split) { trsm_on(st[g], g, k0, k0 + 32, ld, nleft_lo, nleft_hi, c_split); };
    auto gemm_on = [&](cudaStream_t ss, int g, int K, int i0, int kk0, int c0, int c1) {
        // algorithmic columns only (matrix + right-hand side), not the zero padding of ld = Np + 32
        const double mr = Np - i0, nc = (c1 > Np + 1 ? Np + 1 : c1) - c0;
        rt_prof_begin(m, c_gemm, ss, 2. * K * mr * nc, 16. * mr * nc + 8. * K * (mr + nc));
        launch_gemm(d, ss, K, i0, kk0, c0, c1, ga0[g], ga1[g]);
        rt_prof_end(m, c_gemm, ss);
    };
    auto gemm = [&](int g, int K, int i0, int kk0, int c0, int c1) { gemm_on(st[g], g, K, i0, kk0, c0, c1); };
    auto set_layers = [&](int g) { m->prof_layers = layers_all * (ga1[g] - ga0[g]) / std::max(bound, 1); };
    // LOOK-AHEAD schedule (default): per group a second stream T takes everything that is not on the critical path of
    // the next block -- the TRSM of the first panel on the far columns and the trailing update right of the NEXT
    // block's 64 columns -- so the next block's panels (latency-bound, a few SMs) start as soon as their own 64 columns
    // are up to date and run beside the bulk of the trailing update:
    //   S: panel(k0)  trsm [k0+32,k0+64)  gemm32  panel(k0+32)  ..wait T..  trsm2 [k0+64,ld)  gemm64 [k0+64,k0+128)  panel(k0+64) ...
    //   T:            trsm [k0+64,ld)                                                       gemm64 [k0+128,ld)    trsm [k0+128,ld) ...
    // With few active layers (the tail of a solve: every launch is far from filling the GPU) the chain of one block
    // shrinks from panel+trsm+gemm32+panel+trsm+gemm64 to panel+trsm(32 columns)+gemm32+panel+trsm+gemm64(64 columns).
    if (g_lu_lookahead && !m->prof_timing) {
        bool ok = true;
        for (int g = 0; g < G && ok; g++) {
            if (m->lu_side[g]) continue;
            ok = cudaStreamCreateWithPriority(&m->lu_side[g], cudaStreamNonBlocking, m->prio_lo) == cudaSuccess;
            for (int e = 0; e < 3 && ok; e++) ok = cudaEventCreateWithFlags(&m->lu_ev[e][g], cudaEventDisableTiming) == cudaSuccess;
        }
        if (ok) {
            for (int k0 = 0; k0 < Np; k0 += 64) {
                for (int g = 0; g < G; g++) {
                    cudaStream_t S = st[g], T = m->lu_side[g];
                    set_layers(g);
                    panel(g, k0, 0);
                    if (k0 + 32 >= Np) { trsm_on(S, g, k0, k0 + 32, ld, 0, 0, ld); continue; }
                    trsm_on(S, g, k0, k0 + 32, k0 + 64, 0, 0, ld);
                    cudaEventRecord(m->lu_ev[0][g], S);
                    cudaStreamWaitEvent(T, m->lu_ev[0][g], 0);
                    trsm_on(T, g, k0, k0 + 64, ld, 0, 0, ld);
                    cudaEventRecord(m->lu_ev[1][g], T);
                    gemm_on(S, g, 32, k0 + 32, k0, k0 + 32, k0 + 64);
                }
                if (k0 + 32 >= Np) break;
                for (int g = 0; g < G; g++) {
                    cudaStream_t S = st[g], T = m->lu_side[g];
                    set_layers(g);
                    panel(g, k0 + 32, 32);
                    cudaStreamWaitEvent(S, m->lu_ev[1][g], 0);
                    trsm_on(S, g, k0 + 32, k0 + 64, ld, 32, 32, ld);
                    if (k0 + 64 < Np) {
                        const int cmid = std::min(k0 + 128, ld);
                        gemm_on(S, g, 64, k0 + 64, k0, k0 + 64, cmid);
                        if (cmid < ld) {
                            cudaEventRecord(m->lu_ev[2][g], S);
                            cudaStreamWaitEvent(T, m->lu_ev[2][g], 0);
                            gemm_on(T, g, 64, k0 + 64, k0, cmid, ld);
                        }
                    }
                }
            }
            for (int g = 0; g < G; g++) { // the back substitution (on S) needs the far columns (right-hand side) of T
                cudaEventRecord(m->lu_ev[1][g], m->lu_side[g]);
                cudaStreamWaitEvent(st[g], m->lu_ev[1][g], 0);
            }
            goto back_substitution;
        }
    }
    // single stream per group (profiling pass, RT_B200_LU_LOOKAHEAD=0): the groups' launches are issued interleaved, step by
    // step, so that every stream always has work queued
    for (int k0 = 0; k0 < Np; k0 += 64) {
        for (int g = 0; g < G; g++) {
            set_layers(g);
            panel(g, k0, 0);
            trsm(g, k0, 0, 0, ld);
            if (k0 + 32 < Np) gemm(g, 32, k0 + 32, k0, k0 + 32, k0 + 64);
        }
        if (k0 + 32 < Np) {
            for (int g = 0; g < G; g++) {
                set_layers(g);
                panel(g, k0 + 32, 32);
                trsm(g, k0 + 32, 32, 32, ld);
                if (k0 + 64 < Np) gemm(g, 64, k0 + 64, k0, k0 + 64, ld);
            }
        }
    }
back_substitution:
    for (int g = 0; g < G; g++) {
        m->prof_layers = layers_all * (ga1[g] - ga0[g]) / std::max(bound, 1);
        rt_prof_begin(m, c_back, st[g], (double)Np * Np, 4. * Np * Np);
        g_rt_launches++;
        rt_launch(k_backsolve, ga1[g] - ga0[g], BS_THREADS, Np * sizeof(double), st[g], d, ga0[g], ga1[g]);
        rt_prof_end(m, c_back, st[g]);
    }
    m->prof_layers = layers_all;
    for (int g = 1; g < G; g++) {
        cudaEventRecord(m->lu_join[g], st[g]);
        cudaStreamWaitEvent(s0, m->lu_join[g], 0);
    }
}

int lu_max_np() { return PNL_MAXROWS; }

cudaError_t lu_kernels_init()
{
    if (const char* gt = getenv("RT_B200_GEMM32_NARROW")) g_gemm32_narrow = atoi(gt);
    if (const char* gg = getenv("RT_B200_LU_GROUPS")) g_lu_groups = std::max(1, atoi(gg));
    if (const char* gp = getenv("RT_B200_PANEL_192")) g_panel_192 = atoi(gp);
    if (const char* gq = getenv("RT_B200_LU_GROUP_MIN")) g_lu_group_min = std::max(1, atoi(gq));
    if (const char* gl = getenv("RT_B200_LU_LOOKAHEAD")) g_lu_lookahead = atoi(gl);
    if (const char* gd = getenv("RT_B200_PDL_MAX_LAYERS")) g_rt_pdl_max_layers = atoi(gd);
    const int smg = (64 * 36 + 32 * GM_LDB) * (int)sizeof(double);
    const cudaFuncAttribute at = cudaFuncAttributeMaxDynamicSharedMemorySize;
    cudaError_t e = cudaFuncSetAttribute(k_panel<1, 256>, at, 256 * 33 * (int)sizeof(double));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_panel<2, 256>, at, 512 * 33 * (int)sizeof(double));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_panel<2, 192>, at, 384 * 33 * (int)sizeof(double));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_panel<PNL_BIG_R, PNL_BIG_NT>, at, 768 * 33 * (int)sizeof(double));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_trsm<true>, at, (int)trsm_smem_bytes(32));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_gemm<32, 64, 1, 4>, at, smg);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_gemm<32, 64, 2, 4>, at, smg);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_gemm<32, 64, 1, 8, 32>, at, (64 * 36 + 32 * 36) * (int)sizeof(double));
    // ONE preferred shared-memory carve-out (70 % -> 164 KB shared, 92 KB L1) for every LU kernel: CTAs of different kernels
    // (panels, TRSMs, updates of the two layer groups and of the side streams) follow each other on the same SM, and an SM
    // whose resident kernels ask for different carve-outs has to drain before it re-partitions.  Measured per full-batch
    // N = 768 iteration: driver's choice per kernel 4.40 ms; 50 % 4.62; 65-80 % 4.32; 90-100 % 4.87-4.92 (L1 too small for the
    // TRSM gathers); 60 layers 1.95 -> 1.88 ms.  The panels of more than 164 KB get what they need regardless.
    // RT_B200_CARVEOUT=<percent> overrides, -1 leaves the choice to the driver.
    {
        int pct = 70;
        if (const char* gc = getenv("RT_B200_CARVEOUT")) pct = atoi(gc);
        const cudaFuncAttribute ca = cudaFuncAttributePreferredSharedMemoryCarveout;
        if (pct >= 0) {
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_panel<1, 256>, ca, pct);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_panel<2, 256>, ca, pct);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_panel<2, 192>, ca, pct);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_panel<PNL_BIG_R, PNL_BIG_NT>, ca, pct);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_trsm<true>, ca, pct);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_trsm<false>, ca, pct);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_gemm<32, 64, 1, 4>, ca, pct);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_gemm<32, 64, 2, 4>, ca, pct);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_gemm<32, 64, 1, 8, 32>, ca, pct);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_backsolve, ca, pct);
        }
    }
    return e;
}
