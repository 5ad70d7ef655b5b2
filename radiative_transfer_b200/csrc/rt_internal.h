// rt_internal.h -- device-side data layout and kernel launchers of librt_b200.so (not part of the C ABI).
//
// HBM layout (everything FP64, layer-major):
//   Cmat [L][Np][Np]  collisional part of the rate matrix of every layer, in the layout of the reference's
//                     `matrix` (iteration_lvg.cpp:127-131): entry [j][i] = rate i->j, diagonal = -(column sum).
//                     Built ONCE per cloud (it depends on the layer only).
//   Mmat [L][Np][ld]  per-iteration rate matrix = Cmat + radiative terms, row 0 = 1; destroyed by the LU.
//                     ld = Np + 32: column Np carries the right-hand side e0 through the elimination (so the
//                     forward substitution is part of the blocked update), columns Np+1.. are zero padding.
//   Np = N rounded up to a multiple of 32; the padding is an identity block so that every blocked kernel works
//   on whole 32x32 tiles without edge cases (padded unknowns solve to 0 and never couple to real levels).
//   pop  [L][N]       current populations (pop_old of iteration_control::next_step)
//   rhs  [L][Np]      solution of M x = e0 (new populations), written by the back substitution
//   hist_pop/hist_res [L][6][N]  rings of the last nb_prev_steps+1 population / residual vectors
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <cstdint>
#include <string>
#include <vector>
#include "../../include/rt_b200.h"

#define RT_HIST 6
#define RT_MAX_TABLES 16
#define RT_MAX_DUST 8
#define RT_NB 32 // LU panel width == padding granularity
#define RT_LU_MAX_GROUPS 8 // streams one batched LU may be spread over (lu_kernels.cu: launch_lu_solve)
#define RT_AFF_STRIDE 132 // {n_aff, pad, pad, pad, pos[64], cur[64]}

struct DevTable {            // one collision_data, transposed to [jmax][imax] so one layer's T-row is contiguous
    int nb_lev, jmax, imax, interp; // interp: RT_INTERP_* (cubic spline: deriv_t holds the SECOND derivatives)
    double max_temp;
    const double* tgrid;     // [jmax]
    const double* coeff_t;   // [jmax][imax]
    const double* deriv_t;   // [jmax][imax]
};

struct DevLvg { int nb_d, nb_g; const double *delta, *gamma, *p; };
struct DevOverlap { int nb_d, nb_dx, nb_gr, nb_g; const double *log_delta, *dx, *gratio, *gamma, *p; };

// Everything a kernel needs, passed by value (fits the 4 KB parameter space).
struct DevModel {
    int N, Np, R, G, L, ncomp, K, nb1, nb2, rule, scheme, ld; // ld = Np + 32: row stride of Mmat (rhs rides as column Np)
    rt_params prm;
    double mol_mass;
    // levels
    const double* energy; const int* g; const int* v; const double* jq;
    const double* einst;     // [N][N] dense, only read by the boundary-start variant of K0
    // radiative transitions (lines): u > l
    const int *tr_u, *tr_l;
    const double *tr_Aul, *tr_Alu, *tr_E, *tr_sigma; // tr_sigma [R][ncomp] dust cross-section at the line energy
    // groups of 1 or 2 lines whose J is evaluated together: {nb, lineA, lineB}; LVG scheme: G == R, nb == 1
    const int* grp;          // [G][3]
    const double* grp_dxn;   // [G] (E_u1 - E_l1 - E_u2 + E_l2) * c: numerator of dx (iteration_lvg.cpp:455); 0 for singles
    // per-row sparse list of radiative entries: row i holds (col, line, kind) kind 0: i == upper, 1: i == lower
    const int *row_ptr, *row_col, *row_line;
    DevTable tab[RT_MAX_TABLES];
    DevLvg lvg;
    DevOverlap ov[2];
    int ov_same_axes;        // both overlap tables on bitwise identical grids: one index search serves both
    // layers (SoA, [L])
    const double *temp_n, *temp_el, *mol_conc, *vel_width, *vel_grad, *dust_conc /*[L][ncomp]*/;
    const double* conc;      // [L][K] partner concentrations (set_gas_param)
    const int* tidx;         // [L][K] temperature indices (collision_data::locate)
    // big per-layer state
    double *Cmat, *Mmat, *pop, *rhs, *hist_pop, *hist_res, *opt, *yd, *yu;
    int* ipiv;               // [L][Np]
    int* lu_aff;             // [L][Np/32][RT_AFF_STRIDE] row permutation of every panel (see lu_kernels.cu)
    // controller state [L]
    int *iter_nb, *nb_after, *head_pop, *head_res, *n_hist_pop, *n_hist_res, *conv, *active_flag;
    double *best_eq, *eq_err, *pop_err, *rel_err;
    unsigned long long* eq_bits; // [L] max |df| of the current assembly as ordered bits
    int *active, *n_active;  // compacted list of layers still iterating
    int* lay_list;           // [L] scratch: an explicit list of layers handed down by the host (chained start)
};

struct rt_model {
    int device = 0;
    DevModel d{};
    rt_model_desc desc{};          // shallow copy of scalars only
    std::string mol_name;
    std::string last_error;
    std::vector<void*> allocs;     // every cudaMalloc, freed in rt_model_destroy
    std::vector<void*> cloud_allocs;
    std::vector<int> chain_first;  // rt_layer_soa::model_first of the resident cloud ({0, L} for a single slab)
    bool cloud_reuse = false;      // rt_cloud_upload with an unchanged layer count hands the resident buffers out again
    size_t cloud_slot = 0;
    std::vector<std::vector<double>> h_tgrid; // host copy of the collision tables' temperature grids (set_gas_param)
    cudaStream_t stream = nullptr;
    int prio_hi = 0, prio_lo = 0;  // stream priorities: critical chain / far-column bulk work of the look-ahead LU
    cudaStream_t lu_streams[RT_LU_MAX_GROUPS] = {};  // side streams of the grouped LU ([0] unused: group 0 runs on `stream`)
    cudaEvent_t lu_fork = nullptr, lu_join[RT_LU_MAX_GROUPS] = {};
    cudaStream_t lu_side[RT_LU_MAX_GROUPS] = {};     // look-ahead: far-column work of every group
    cudaEvent_t lu_ev[3][RT_LU_MAX_GROUPS] = {};
    int* h_n_active = nullptr;     // pinned
    bool has_cloud = false;
    bool use_fused = true;         // persistent one-CTA-per-layer solver when the matrix fits in shared memory
    // host copies needed by host-side entry points
    std::vector<double> h_energy, h_einst;
    std::vector<int> h_tr_u, h_tr_l, h_grp;
    std::vector<rt_layer_status> h_status;
    std::vector<double> h_dz, h_vel_n, h_vel_turb;
    std::vector<double> h_h2_conc, h_mol_conc, h_velg; // rt_lim_luminosity: ph2 + oh2, mol_conc, dv/dz per layer
    std::vector<int> h_g;
    double h_height = 0.;
    int gw_up = -1, gw_low = -1;   // rt_set_gain_width_line
    double gw_offset = 0.;
    long long iters_enqueued = 0;
    std::vector<std::vector<double>> h_dust_en, h_dust_abs;
    std::vector<double> h_dust_exp;
    // profiling
    bool prof_timing = false;
    std::vector<std::string> prof_names;
    std::vector<double> prof_ms;
    std::vector<long long> prof_launches;
    std::vector<double> prof_flops, prof_bytes; // algorithmic work of the launches, summed over the layers they process
    struct ProfPending { int cls; cudaEvent_t a, b; };
    std::vector<ProfPending> prof_pending;
    std::vector<cudaEvent_t> ev_pool;
    int prof_layers = 0;           // layers the launches enqueued now process (set by the drivers in rt_api.cu)
    cudaEvent_t timer_ev[2] = {nullptr, nullptr};
};

extern std::atomic<long long> g_rt_launches; // every kernel launch of the library (bench.py gpu_launches); several host
                                             // threads may drive their own rt_model handles concurrently

// Programmatic dependent launch for the kernels of the per-iteration chain.  An iteration with few active layers is a chain
// of ~75 dependent kernels of 5-30 us at their latency floor; with the attribute below kernel k+1 is launched, its CTAs
// scheduled and its preamble run while kernel k drains.  Every kernel launched through rt_launch begins with rt_pdl_sync():
// `griddepcontrol.wait` returns when the preceding grid has completed and its writes are visible (a no-op without the
// attribute), `griddepcontrol.launch_dependents` lets the next kernel of the stream start its own launch.  Waiting CTAs hold
// SM slots, which costs more than it saves when another stream could use them (the grouped LU of a full batch), so the
// driver switches it on while active layers x Np^2 <= g_rt_pdl_max_layers x 768^2 (RT_B200_PDL_MAX_LAYERS, default 96, 0 =
// never).  Measured per iteration, N = 768: 8 layers 1.24 -> 1.13 ms, 25 layers 1.41 -> 1.34 ms, 60 layers equal, 200
// layers 4.46 -> 4.58 ms (hence the threshold); o-H2O N = 411, 200 layers: converged solve 92.7 -> 87.8 ms.  Results are
// bitwise unchanged (only launch timing moves).
extern thread_local int g_rt_pdl_on;
extern int g_rt_pdl_max_layers;
#ifdef __CUDACC__
static __device__ __forceinline__ void rt_pdl_sync()
{
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
template <class... KArgs, class... Args>
static inline void rt_launch(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = g_rt_pdl_on ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}
#endif

// ---- kernel launchers (each returns after enqueueing on `s`) -------------------------------------------------
// layers processed = active[0 .. min(*n_active, nlay_bound))
void launch_coll_matrix(const DevModel& d, cudaStream_t s, double* out, int neutrals_only_half_A, int active_bound = 0);
void launch_intensity(const DevModel& d, cudaStream_t s, int nlay_bound);
void launch_assemble(const DevModel& d, cudaStream_t s, int nlay_bound, double* df_out /*[L][N] or null*/);
void launch_lu_solve(rt_model* m, cudaStream_t s, int nlay_bound);
void launch_ctl_init_flagged(const DevModel& d, cudaStream_t s, int accel);
void launch_ctl_step(const DevModel& d, cudaStream_t s, int nlay_bound, int accel, int max_iter, double min_err);
void launch_compact(const DevModel& d, cudaStream_t s);
void launch_set_active_range(const DevModel& d, cudaStream_t s, int lay_begin, int lay_end);
void launch_set_active_failed(const DevModel& d, cudaStream_t s, int lay_begin, int lay_end);
void launch_set_active_list(const DevModel& d, cudaStream_t s, int n);   // flags <- d.lay_list[0..n)
void launch_copy_prev_pop(const DevModel& d, cudaStream_t s, int n);     // pop[lay] <- pop[lay-1] for lay in d.lay_list[0..n)
void launch_boltzmann(const DevModel& d, cudaStream_t s);
void launch_set_rhs(const DevModel& d, cudaStream_t s, int nlay_bound);
void launch_set_row0(const DevModel& d, cudaStream_t s, int nlay_bound);
void launch_copy_rhs_to_pop(const DevModel& d, cudaStream_t s, int nlay_bound);
void launch_intensity_single(const DevModel& d, cudaStream_t s, int lay, int up, int low, double Aul, double Alu,
                             int pair_up, int pair_low, double pAul, double pAlu, int which, const double* pop_dev,
                             double* out_dev);

// profiling hooks used by the launchers
int rt_prof_class(rt_model* m, const char* name);
// flops / bytes: ALGORITHMIC work per layer of this launch (DESIGN.md section "Kernels"); multiplied by m->prof_layers
void rt_prof_begin(rt_model* m, int cls, cudaStream_t s, double flops_per_layer = 0., double bytes_per_layer = 0.);
void rt_prof_end(rt_model* m, int cls, cudaStream_t s);
