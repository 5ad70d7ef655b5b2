/*
 * rt_b200.h -- C ABI of the B200-native level-population solver (librt_b200.so).
 *
 * This is the drop-in boundary for the hot path of nesterenok/radiative_transfer:
 *     calc_molecular_populations()                     radiative_transfer.cpp:219-289
 *       -> iteration_control<T>::calculate_populations iteration_control.h:196-242
 *         -> iteration_scheme_lvg::calc_new_pop        iteration_lvg.cpp:87-110
 *           -> operator() / intensity_calc / lu_matrix_solve
 *
 * The reference has no FFI: the seam is a C++ call.  A maintainer flattens the reference's objects
 * (energy_diagram, einstein_coeff, collisional_transitions, lvg_method_data, dust_model, cloud_data) into the
 * plain structs below ONCE per model and calls rt_solve(); see INTEGRATION.md for the adapter
 * (`calc_molecular_populations_gpu`, same signature as radiative_transfer.cpp:219).
 *
 * Everything is FP64.  No torch types, no C++ types: plain pointers and sizes.  All pointers in the *_desc
 * structs are HOST pointers that only need to stay valid during the call they are passed to.
 * Fatal conditions that make the reference print and exit(1) are returned as status codes.
 * There is NO CPU fallback: every entry point that computes returns RT_ERR_CUDA if no sm_100 device is usable.
 */
#ifndef RT_B200_H
#define RT_B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef enum rt_status {
    RT_OK = 0,
    RT_ERR_INVALID = 1,           /* bad argument / inconsistent sizes                                    */
    RT_ERR_CUDA = 2,              /* CUDA runtime error or no usable device (see rt_last_error)           */
    RT_ERR_NOMEM = 3,
    RT_ERR_DUST_INCONSISTENT = 4, /* dust vectors != dust->nb_of_comp: iteration_lvg.cpp:81-84 (exit(1))  */
    RT_ERR_NO_CLOUD = 5,          /* rt_solve before rt_cloud_upload                                      */
    RT_ERR_UNSUPPORTED = 6
} rt_status;

/* Species mixing rule of collisional_transitions::get_rate_neutrals */
enum {
    RT_RULE_GENERIC = 0, /* collisional_transitions::get_rate_neutrals        coll_rates.cpp:181-197        */
    RT_RULE_CH3OH = 1,   /* ch3oh_collisions::get_rate_neutrals               coll_rates_ch3oh.cpp:484-533  */
    RT_RULE_OH_HF = 2,   /* oh_hf_collisions::get_rate_neutrals               coll_rates_oh.cpp:392-407     */
    RT_RULE_H2O = 3,     /* h2o_collisions::get_rate_neutrals                 coll_rates_h2o.cpp:530-548    */
    RT_RULE_NH3 = 4,     /* nh3_collisions::get_rate_neutrals (He, p-H2, o-H2, H) coll_rates_nh3.cpp:261-290     */
    RT_RULE_H2CO = 5     /* h2co_collisions::get_rate_neutrals (He, p-H2, o-H2) coll_rates_h2co.cpp:160-185      */
};

/* Iteration scheme */
enum {
    RT_SCHEME_LVG = 0,         /* iteration_scheme_lvg            iteration_lvg.h:17-65   */
    RT_SCHEME_LINE_OVERLAP = 1 /* iteration_scheme_line_overlap   iteration_lvg.h:87-106  */
};

/* Start vector of every layer */
enum {
    RT_START_GIVEN = 0,    /* pop[] holds the start vector of every layer                                   */
    RT_START_BOUNDARY = 1, /* boundary_layer_populations() for every layer   iteration_control.cpp:52-91    */
    RT_START_LTE = 2,      /* boltzmann_populations(T_n)                     spectroscopy.cpp:1245-1259     */
    RT_START_CHAINED = 3   /* reference warm-start chain                     radiative_transfer.cpp:247-253 */
};

/* Compile-time constants of the reference, carried as data (SURVEY.md section 5 "Config / flags"). */
typedef struct rt_params {
    double inv_trans_factor;      /* INV_TRANS_FACTOR  -0.1      iteration_lvg.cpp:23   */
    double min_line_opacity;      /* MIN_LINE_OPACITY  1e-99     iteration_lvg.cpp:24   */
    double min_collision_rate;    /* MIN_COLLISION_RATE 1e-99    coll_rates.h:9         */
    double max_dx;                /* max_dx = 4                  iteration_lvg.cpp:307  */
    int accel_start;              /* 40                          iteration_control.h:71 */
    int accel_period;             /* 5                           iteration_control.h:71 */
    int nb_prev_steps;            /* 5 (only 5 is implemented)   iteration_control.h:71 */
    int reserved0;
    double boltzmann_constant;    /* erg/K     constants.h (missing from the reference; oracle/shim) */
    double cm_inverse_to_kelvins; /* hc/k                                                            */
    double speed_of_light;        /* cm/s                                                            */
    double eight_pi;
    double onedivby_sqrt_pi;
    double planck_constant;       /* erg s (ABI 2; rt_lim_luminosity: temp_sat, transition_data.cpp:479) */
} rt_params;

/* One collision_data table (coll_rates.h:12-41).  coeff is [imax][jmax] row-major, imax = nb_lev*(nb_lev-1)/2,
 * row index i*(i-1)/2 + j for the transition i -> j (i > j).  The library derives coeff_deriv itself:
 * RT_INTERP_LINEAR        collision_data: first differences (calc_coeff_deriv, coll_rates.cpp:61-69), linear get_rate (:54-59);
 * RT_INTERP_CUBIC_SPLINE  collision_data_cub_spline (coll_rates.h:44-53): natural-spline second derivatives
 *                         (calc_coeff_deriv, coll_rates.cpp:89-119) and the cubic get_rate with its fabs (:121-132).  ABI 5. */
enum { RT_INTERP_LINEAR = 0, RT_INTERP_CUBIC_SPLINE = 1 };
typedef struct rt_coll_table {
    int nb_lev;
    int jmax;
    const double* tgrid; /* [jmax] ascending */
    const double* coeff; /* [imax][jmax]     */
    int interp;          /* RT_INTERP_* */
    int reserved0;
} rt_coll_table;

/* One dust_component as seen by dust_component::absorption (dust_model.cpp:473-490). */
typedef struct rt_dust_comp {
    int nb_ph_en;
    int reserved0;
    const double* ph_en;     /* [nb_ph_en] ascending photon energies, cm-1  */
    const double* abs_coeff; /* [nb_ph_en] absorption cross-section per grain, cm2 */
    double wvl_exp;
} rt_dust_comp;

/* lvg_method_data (lvg_method_functions.h:23-40): p[nb_d][nb_g], delta first index. */
typedef struct rt_lvg_table {
    int nb_d, nb_g;
    const double* delta; /* [nb_d] */
    const double* gamma; /* [nb_g] */
    const double* p;     /* [nb_d][nb_g] */
} rt_lvg_table;

/* lvg_line_overlap_data (lvg_method_functions.h:53-69): p[(d*nb_dx + dx)][(gr*nb_g + g)].
 * delta is given LINEAR (as in the data file); the library takes log10 as lvg_method_functions.cpp:293. */
typedef struct rt_overlap_table {
    int nb_d, nb_dx, nb_gr, nb_g;
    const double* delta;  /* [nb_d]  */
    const double* dx;     /* [nb_dx] */
    const double* gratio; /* [nb_gr] */
    const double* gamma;  /* [nb_g]  */
    const double* p;      /* [nb_d*nb_dx][nb_gr*nb_g] */
    int delta_is_log10;   /* != 0: delta[] already holds log10(delta), as lvg_line_overlap_data keeps it in memory */
    int reserved0;
} rt_overlap_table;

/* Molecule + physics tables: replaces energy_diagram (spectroscopy.h:69-87), einstein_coeff (:179-190),
 * collisional_transitions (coll_rates.h:56-99), lvg_method_data / lvg_line_overlap_data, dust_model. */
typedef struct rt_model_desc {
    int nb_lev;           /* N = energy_diagram::nb_lev                                                    */
    int species_rule;     /* RT_RULE_*                                                                     */
    const double* energy; /* [N] cm-1, energy_level::energy (ascending as in the reference's diagrams)     */
    const int* g;         /* [N] statistical weights, energy_level::g                                      */
    const int* v;         /* [N] vibrational / torsional quantum number (CH3OH rule), energy_level::v      */
    const double* j;      /* [N] angular momentum (CH3OH rule: J <= 9 test), energy_level::j               */
    double mol_mass;      /* g, molecule::mass                                                             */
    const char* mol_name; /* molecule::name ("CH3OHa"/"CH3OHe" suppress the un-accelerated retry)          */
    const double* einst;  /* [N][N] row-major == einstein_coeff::arr (arr[u][l]=A_ul, arr[l][u]=A_ul g_u/g_l) */

    int nb_neutral_tables;  /* nb1: tables [0, nb1) are neutral partners                                   */
    int nb_electron_tables; /* nb2 - nb1: tables [nb1, nb2) are electron data (first applicable is used)   */
    const rt_coll_table* tables;

    int scheme;             /* RT_SCHEME_*                                                                 */
    int nb_line_groups;     /* RT_SCHEME_LINE_OVERLAP: number of hfs groups, 0 = build them in the library
                               with the rule of iteration_scheme_line_overlap::init_molecule_data
                               (iteration_lvg.cpp:317-346)                                                 */
    const int* line_groups; /* [nb_line_groups][5] = {nb(1|2), up0, low0, up1, low1}                       */
    rt_lvg_table lvg;       /* lvg/lvg_loss_func.txt                                                       */
    const rt_overlap_table* overlap1; /* lvg/line_overlap_func_p1.txt (RT_SCHEME_LINE_OVERLAP only)        */
    const rt_overlap_table* overlap2; /* lvg/line_overlap_func_p2.txt                                      */

    int dust_ncomp;           /* dust_model::nb_of_comp                                                    */
    int reserved0;
    const rt_dust_comp* dust; /* [dust_ncomp]                                                              */

    rt_params params;
} rt_model_desc;

/* Cloud: replaces cloud_data / cloud_layer (cloud_data.h:24-78) as a layer-major SoA. */
typedef struct rt_layer_soa {
    int nb_lay;        /* L */
    int dust_ncomp;    /* must equal rt_model_desc::dust_ncomp, else RT_ERR_DUST_INCONSISTENT */
    const double* temp_n;   /* [L] each */
    const double* temp_el;
    const double* el_conc;
    const double* h_conc;
    const double* ph2_conc;
    const double* oh2_conc;
    const double* he_conc;
    const double* mol_conc;
    const double* vel_turb;
    const double* velg_n;   /* dv/dz, signed */
    const double* dz;       /* post-pass only (transition_data.cpp) */
    const double* vel_n;    /* post-pass only */
    const double* dust_conc; /* [L][dust_ncomp] */
    double height;          /* cloud_data::get_height(), post-pass only */
    /* Optional (ABI 2): the state an iteration_scheme_lvg object holds after its setters ran, passed through verbatim so a
     * scheme subclass reproduces it bit for bit.  NULL = derived from the fields above. */
    const double* vel_width;    /* [L] iteration_lvg.cpp:65; NULL: sqrt(2kT/m + vel_turb^2) */
    const double* partner_conc; /* [L][nb_neutral_tables + nb_electron_tables] coll_partn_conc of set_gas_param; NULL: species rule */
    /* Optional (ABI 4): the cloud is a STACK of nb_models independent slabs of the same molecule (a parameter sweep solved
     * as one batch): model k owns layers [model_first[k], model_first[k+1]).  Layers are independent, so this changes
     * nothing for the independent-start modes; RT_START_CHAINED then runs the reference's warm-start chain
     * (radiative_transfer.cpp:247-253) inside every model and advances layer l of ALL models together, one nb_models-wide
     * batch per chain step, never chaining across a model boundary.  nb_models == 0 / NULL: one model of nb_lay layers. */
    int nb_models;
    int reserved0;
    const int* model_first;     /* [nb_models + 1], model_first[0] == 0, ascending, model_first[nb_models] == nb_lay */
} rt_layer_soa;

/* Per-layer outcome == the public members of iteration_control after calculate_populations
 * (iteration_control.h:54-57) plus the iteration count. */
typedef struct rt_layer_status {
    int iters;     /* iter_nb on exit */
    int converged; /* return value of calculate_populations */
    double eq_error, pop_error, rel_error;
} rt_layer_status;

/* Opaque handle: one molecule's tables + one resident cloud on one GPU.  No global state in the library: a handle is used by
 * one host thread at a time; DIFFERENT handles (also on the same GPU) may be driven concurrently from different host threads
 * -- bench.py's sweep keeps three models in flight per GPU that way (each handle owns its streams and buffers; host-to-device
 * uploads are ordered on the handle's own stream). */
typedef struct rt_model rt_model;

void rt_params_default(rt_params* p);
const char* rt_status_string(rt_status s);
const char* rt_last_error(const rt_model* m);
int rt_abi_version(void);

/* Flatten + upload the model to `device` (CUDA ordinal).  Builds transposed collision tables, transition
 * records, dust cross-sections per transition (dust_component::absorption, dust_model.cpp:473-490). */
rt_status rt_model_create(const rt_model_desc* desc, int device, rt_model** out);
rt_status rt_model_destroy(rt_model* m);

/* == the three setters called per layer in radiative_transfer.cpp:238-242, for all layers at once; also builds
 * the per-layer collisional matrix C(layer) on device (set_gas_param + get_rate_neutrals/electrons). */
rt_status rt_cloud_upload(rt_model* m, const rt_layer_soa* layers);

/* Re-runs the collisional-matrix kernel from the layer arrays already resident in HBM (what rt_cloud_upload does
 * after its host-to-device copies); bench.py's device-resident `value` leg calls it so that every timed step
 * contains the whole per-cloud work. */
rt_status rt_cloud_rebuild(rt_model* m);

/* == boundary_layer_populations for every layer (iteration_control.cpp:52-91); pop is HOST [L*N]. */
rt_status rt_boundary_populations(rt_model* m, double* pop);

/* == iteration_scheme_lvg::calc_new_pop (iteration_lvg.cpp:87-110) for every layer: one step, no controller.
 * old_pop/new_pop HOST [L*N]; eq_err HOST [L]. */
rt_status rt_step(rt_model* m, const double* old_pop, double* new_pop, double* eq_err);

/* == iteration_scheme_lvg::operator() (iteration_lvg.cpp:112-161): returns the assembled matrix (row 0 = 1)
 * and df for every layer.  matrix HOST [L][N][N] (may be NULL), df HOST [L][N] (may be NULL). */
rt_status rt_assemble(rt_model* m, const double* pop, double* matrix, double* df);

/* == intensity_calc(up, low, level_pop, J) for layer `lay` (iteration_lvg.cpp:163-185; with the overlap scheme
 * the single-line overload :503-526).  pop HOST [N] for that layer. */
rt_status rt_intensity(rt_model* m, int lay, int up, int low, const double* pop, double* intensity);

/* rt_intensity without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments and result as rt_intensity; the device copies of `pop` and of the result block live in a file-scope
 * device array of RT_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit, so nb_lev is limited to
 * RT_NOALLOC_MAX_LEVELS (RT_ERR_UNSUPPORTED beyond, message in rt_last_error).  Calls no cudaMalloc / cudaFree /
 * allocating routine.  All calls on one GPU share the array: calls that overlap on the GPU (other host threads, other
 * rt_model handles on the same device) race -- serialise them.  The call itself is synchronous. */
#define RT_NOALLOC_MAX_LEVELS 1024
#define RT_NOALLOC_SCRATCH_BYTES ((RT_NOALLOC_MAX_LEVELS + 16) * 8)
rt_status rt_intensity_noalloc(rt_model* m, int lay, int up, int low, const double* pop, double* intensity);

/* == iteration_control::calculate_populations for every layer (iteration_control.h:196-242), all layers at
 * once on the device.  pop HOST [L*N] in/out (input used when start_mode == RT_START_GIVEN). */
rt_status rt_solve(rt_model* m, double* pop, int start_mode, int max_iter, double min_err, int accel,
                   rt_layer_status* st);

/* == calc_molecular_populations (radiative_transfer.cpp:219-289) incl. the retry policy (:259-276):
 * accelerated run with max_iter_acc, then for layers that failed and molecules other than CH3OHa/CH3OHe an
 * un-accelerated run with max_iter_ext.  start_mode RT_START_CHAINED reproduces the warm-start chain (one chain per
 * model of a stacked cloud, see rt_layer_soa::model_first; chain step l of every model is one batch on the device). */
rt_status rt_calc_molecular_populations(rt_model* m, double* pop, int start_mode, int acceleration,
                                        int max_iter_acc, int max_iter_ext, double min_err, rt_layer_status* st);

/* Device-resident variant used by bench.py `value`: populations stay in HBM (rt_device_pop), no H2D/D2H in the
 * call; status is fetched with rt_get_status. */
rt_status rt_solve_device(rt_model* m, int start_mode, int max_iter, double min_err, int accel);
rt_status rt_get_status(rt_model* m, rt_layer_status* st);
rt_status rt_get_populations(rt_model* m, double* pop);
rt_status rt_set_populations(rt_model* m, const double* pop);
/* Populations [L*N] copied device-to-device into a caller-owned DEVICE buffer on the model's GPU (the sharded solve hands
 * them to the NCCL all_gather without a host round trip).  Synchronous: complete on return. */
rt_status rt_copy_populations_device(rt_model* m, double* dst_device);

/* Instrumentation for bench.py / profiles: kernel launch count and CUDA-event time (ms) per kernel class since
 * the last reset.  names/ms/launches arrays of capacity `cap`; returns number of classes through *n. */
rt_status rt_profile_reset(rt_model* m, int enable_timing);
rt_status rt_profile_get(rt_model* m, int cap, const char** names, double* ms, long long* launches, int* n);
long long rt_iterations_enqueued(const rt_model* m); /* controller steps enqueued since rt_profile_reset */
long long rt_launch_count(void); /* kernels launched by this library in this process so far (bench.py gpu_launches) */
/* ALGORITHMIC flops / bytes of the launches of each class since the last reset (same order as rt_profile_get):
 * per-layer figure of DESIGN.md x the layers each launch processed. */
rt_status rt_profile_get_work(rt_model* m, int cap, double* flops, double* bytes);
/* CUDA-event stopwatch on the library's own stream (torch.cuda.Event only sees torch's stream). */
rt_status rt_timer_begin(rt_model* m);
rt_status rt_timer_end(rt_model* m, double* ms);
/* Measured FP64 ceilings of `device`: a register-resident DFMA loop and a DMMA (mma.sync.m8n8k4.f64) loop on every
 * SM, TFLOP/s, best of `reps`.  MEASURED_PEAKS.json holds no FP64 figure, so bench.py measures the denominator of
 * the LU roofline in the same run. */
rt_status rt_device_fp64_peak(int device, int reps, double* dfma_tflops, double* dmma_tflops);

/* Post-solve maser quantities (transition_data.cpp:210-378): for each listed transition (up[i] > low[i]):
 * inv[L], gain[L] per layer and the scalars {inv, gain, tau_eff, tau_max, lay_nb_hg}; tau_vs_aspect [na],
 * tau_vs_freq [nf].  pop HOST [L*N]. Arrays may be NULL. */
rt_status rt_maser_postpass(rt_model* m, const double* pop, int nb_trans, const int* up, const int* low,
                            int nb_freq, int nb_aspect, double delta_aspect, double velocity_shift,
                            double* inv_arr, double* gain_arr, double* scalars /*[nb_trans][5]*/,
                            double* tau_vs_aspect, double* tau_vs_freq);

/* The oH2O 6_16 -> 5_23 (22.2 GHz) special case of calc_gain (transition_data.cpp:249-260): the gain of that one line
 * uses vel = sqrt((sqrt(2kT/m) + thermal_offset)^2 + v_turb^2), thermal_offset = 5e4 cm/s in the reference (hyperfine
 * splitting).  The reference recognises the line by the quantum numbers (v, j, k1, k2) of its levels, which the flat
 * model does not carry: the caller (who holds the energy_diagram) names it by level indices once per model.
 * up < 0 clears it.  Affects rt_maser_postpass only (gain, tau_eff, lay_nb_hg; not the line profile, as in the
 * reference).  ABI 3. */
rt_status rt_set_gain_width_line(rt_model* m, int up, int low, double thermal_offset);

/* Limiting maser luminosity and pump/loss rates: lim_luminosity_lvg (maser_luminosity.cpp:7-106),
 * calc_exc_temp (transition_data.cpp:225-237) and calc_saturation_depth (transition_data.cpp:456-483) for each listed
 * transition (up[i] > low[i]).  pop HOST [L*N].  Per transition and layer, all [nb_trans][L], any may be NULL:
 * lum_arr, pump_eff_arr, loss_rate_arr, pump_rate_arr, emiss_coeff_arr, exc_temp_arr; per transition: lum [nb_trans].
 * layer0_intensity != 0 reproduces the reference literally: the mean intensity inside the radiative loss terms is
 * evaluated with the populations of LAYER 0 and the gas parameters of the current layer (intensity_calc is handed the
 * base pointer `level_pop`, maser_luminosity.cpp:54,58); 0 uses the populations of the current layer.
 * Saturation: with lay_nb_hg [nb_trans] (layer of the highest gain, scalars[4] of rt_maser_postpass) and a beaming
 * factor, tau_sat / temp_sat [nb_trans] are filled for transitions whose excitation temperature in that layer is
 * negative and left at 0 otherwise; lay_nb_hg == NULL skips it.
 * Needs the cloud uploaded with the plain rt_layer_soa fields (ph2_conc, oh2_conc, dz, height). */
rt_status rt_lim_luminosity(rt_model* m, const double* pop, int nb_trans, const int* up, const int* low,
                            int layer0_intensity, double* lum_arr, double* pump_eff_arr, double* loss_rate_arr,
                            double* pump_rate_arr, double* emiss_coeff_arr, double* exc_temp_arr, double* lum,
                            const int* lay_nb_hg, double beaming_factor, double* tau_sat, double* temp_sat);

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_H */
