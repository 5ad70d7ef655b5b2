// rt_api.cu -- the extern "C" layer of librt_b200.so (include/rt_b200.h): flattens the host description into the
// device layout of rt_internal.h, owns all device memory, and drives the per-iteration kernel sequence.
// No CPU fallback: every computing entry point needs a CUDA device and returns RT_ERR_CUDA otherwise.
#include "rt_internal.h"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cstdlib>

cudaError_t lu_kernels_init();
cudaError_t fused_kernels_init();
bool fused_supported(const DevModel& d);
void launch_fused_solve(const DevModel& d, cudaStream_t s, int nlay, int accel, int max_iter, double min_err);
int lu_max_np();
void launch_ctl_init_flagged(const DevModel& d, cudaStream_t s, int accel);
void launch_set_active_range(const DevModel& d, cudaStream_t s, int b, int e);
void launch_set_active_failed(const DevModel& d, cudaStream_t s, int b, int e);

#define CK(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess) {                                                                         \
            char buf_[512];                                                                              \
            snprintf(buf_, sizeof buf_, "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            m->last_error = buf_;                                                                        \
            return RT_ERR_CUDA;                                                                          \
        }                                                                                                \
    } while (0)

static thread_local std::string g_create_error;
std::atomic<long long> g_rt_launches{0};
thread_local int g_rt_pdl_on = 0;
int g_rt_pdl_max_layers = 96;

template <class T> static T* dev_alloc(rt_model* m, size_t n, bool cloud, bool zero = true)
{
    void* p = nullptr;
    if (n == 0) n = 1;
    if (cloud && m->cloud_reuse) {
        // same layer count as the resident cloud: every buffer has the size it had, in the same order; a parameter sweep
        // then never touches the allocator (cudaMalloc/cudaFree serialise across the GPUs of a node)
        p = m->cloud_allocs[m->cloud_slot++];
        if (zero) cudaMemsetAsync(p, 0, n * sizeof(T), m->stream);
        return (T*)p;
    }
    if (cudaMalloc(&p, n * sizeof(T)) != cudaSuccess) { m->last_error = "cudaMalloc failed"; return nullptr; }
    if (zero) cudaMemsetAsync(p, 0, n * sizeof(T), m->stream);
    (cloud ? m->cloud_allocs : m->allocs).push_back(p);
    return (T*)p;
}

template <class T> static T* dev_upload(rt_model* m, const T* h, size_t n, bool cloud)
{
    T* p = dev_alloc<T>(m, n, cloud, false);
    if (p && n) {
        // ON THE MODEL'S STREAM: a synchronous cudaMemcpy from pageable host memory returns once the data are staged -- the
        // DMA to the device may still be in flight on the legacy stream, which a non-blocking stream does not wait for, so a
        // kernel enqueued next on m->stream could read the previous cloud's values (seen as run-to-run differences of a sweep
        // with several models in flight).  Stream order makes the kernels of m->stream wait for the copy; the host buffer is
        // free again on return (pageable: staged immediately; pinned: the callers synchronise before the buffer goes away).
        if (cudaMemcpyAsync(p, h, n * sizeof(T), cudaMemcpyHostToDevice, m->stream) != cudaSuccess) { m->last_error = "cudaMemcpy H2D failed"; return nullptr; }
    }
    return p;
}

// scratch device buffers of one entry-point call: freed on every exit path (the CK macro returns early)
struct DevScratch {
    std::vector<void*> bufs;
    cudaStream_t stream; // uploads are ordered on the stream whose kernels read them (see dev_upload)
    explicit DevScratch(cudaStream_t s) : stream(s) {}
    ~DevScratch() { for (void* p : bufs) cudaFree(p); }
    void* get(size_t bytes, const void* host = nullptr)
    {
        void* p = nullptr;
        if (cudaMalloc(&p, bytes ? bytes : 1) != cudaSuccess) return nullptr;
        bufs.push_back(p);
        if (host && cudaMemcpyAsync(p, host, bytes, cudaMemcpyHostToDevice, stream) != cudaSuccess) return nullptr;
        return p;
    }
};
#define NEED_BUF(ptr) do { if (!(ptr)) { m->last_error = "device scratch allocation / upload failed"; return RT_ERR_NOMEM; } } while (0)

extern "C" {

void rt_params_default(rt_params* p)
{
    if (!p) return;
    memset(p, 0, sizeof *p);
    p->inv_trans_factor = -0.1;
    p->min_line_opacity = 1.e-99;
    p->min_collision_rate = 1.e-99;
    p->max_dx = 4.;
    p->accel_start = 40;
    p->accel_period = 5;
    p->nb_prev_steps = 5;
    p->boltzmann_constant = 1.38064852e-16;
    p->cm_inverse_to_kelvins = 1.438777364;
    p->speed_of_light = 2.99792458e+10;
    p->eight_pi = 25.132741228718345;
    p->onedivby_sqrt_pi = 0.56418958354775629;
    p->planck_constant = 6.62607004e-27;
}

const char* rt_status_string(rt_status s)
{
    switch (s) {
    case RT_OK: return "ok";
    case RT_ERR_INVALID: return "invalid argument";
    case RT_ERR_CUDA: return "CUDA error / no device";
    case RT_ERR_NOMEM: return "out of memory";
    case RT_ERR_DUST_INCONSISTENT: return "inconsistency in the interstellar dust data";
    case RT_ERR_NO_CLOUD: return "no cloud uploaded";
    case RT_ERR_UNSUPPORTED: return "unsupported configuration";
    }
    return "unknown";
}

const char* rt_last_error(const rt_model* m) { return m ? m->last_error.c_str() : g_create_error.c_str(); }
int rt_abi_version(void) { return 5; }

} // extern "C"

// dust_component::absorption(energy), dust_model.cpp:473-490 -- evaluated on the host once per line
static double host_dust_sigma(const rt_dust_comp& c, double energy)
{
    const int n = c.nb_ph_en;
    if (energy < c.ph_en[0]) return c.abs_coeff[0] * pow(c.ph_en[0] / energy, c.wvl_exp);
    if (energy > c.ph_en[n - 1]) return c.abs_coeff[n - 1];
    int l = 0, r = n - 1;
    while (r - l > 1) {
        const int i = l + ((r - l) >> 1);
        if (c.ph_en[i] < energy) l = i; else r = i;
    }
    // abs_coeff_deriv of dust_component::calc_derivatives, dust_model.cpp:516-524
    const double deriv = (c.abs_coeff[l + 1] - c.abs_coeff[l]) / (c.ph_en[l + 1] - c.ph_en[l]);
    return c.abs_coeff[l] + deriv * (energy - c.ph_en[l]);
}

// hfs_lines::sort / split + iteration_scheme_line_overlap::init_molecule_data, iteration_lvg.cpp:259-346
namespace {
struct Hfs { int nb = 0; int up[4], low[4]; double en[4]; };
void hfs_swap(Hfs& h, int i, int j) { std::swap(h.up[i], h.up[j]); std::swap(h.low[i], h.low[j]); std::swap(h.en[i], h.en[j]); }
void hfs_sort(Hfs& h)
{
    if (h.nb <= 2) return;
    for (int i = 0; i < h.nb; i++)
        for (int j = i + 1; j < h.nb; j++)
            if (h.en[j] < h.en[i]) hfs_swap(h, i, j);
    int j = 0;
    double e = h.en[1] - h.en[0];
    for (int i = 1; i < h.nb - 1; i++)
        if (h.en[i + 1] - h.en[i] < e) j = i; // `e` is never updated in the reference (:274-280): kept
    e = h.en[j];
    for (int i = 0; i < h.nb; i++)
        for (int jj = i + 1; jj < h.nb; jj++)
            if (fabs(h.en[jj] - e) < fabs(h.en[i] - e)) hfs_swap(h, i, jj);
}
std::vector<int> build_line_groups(const rt_model_desc* d)
{
    std::vector<int> g;
    const int N = d->nb_lev;
    for (int i = 2; i < N; i += 2)
        for (int j = 0; j < i; j += 2) {
            Hfs h;
            for (int mm = 0; mm < 2; mm++)
                for (int l = 0; l < 2; l++)
                    if (i + mm < N && d->einst[(size_t)(i + mm) * N + j + l] > 1.e-99) {
                        h.up[h.nb] = i + mm; h.low[h.nb] = j + l;
                        h.en[h.nb] = d->energy[i + mm] - d->energy[j + l];
                        h.nb++;
                    }
            hfs_sort(h);
            int first = 0;
            if (h.nb >= 3) { g.insert(g.end(), {2, h.up[0], h.low[0], h.up[1], h.low[1]}); first = 2; }
            const int rest = h.nb - first;
            if (rest > 0)
                g.insert(g.end(), {rest, h.up[first], h.low[first], rest > 1 ? h.up[first + 1] : -1, rest > 1 ? h.low[first + 1] : -1});
        }
    return g;
}
} // namespace

extern "C" rt_status rt_model_create(const rt_model_desc* desc, int device, rt_model** out)
{
    if (!desc || !out) { g_create_error = "null argument"; return RT_ERR_INVALID; }
    *out = nullptr;
    const int N = desc->nb_lev;
    const int K = desc->nb_neutral_tables + desc->nb_electron_tables;
    if (N < 2 || !desc->energy || !desc->g || !desc->einst || K > RT_MAX_TABLES || desc->dust_ncomp > RT_MAX_DUST
        || desc->dust_ncomp < 1 || !desc->lvg.p) { g_create_error = "inconsistent rt_model_desc"; return RT_ERR_INVALID; }
    if ((N + RT_NB - 1) / RT_NB * RT_NB > lu_max_np()) { g_create_error = "nb_lev too large for the register-resident LU panel (max 768, the reference's largest molecule)"; return RT_ERR_UNSUPPORTED; }
    if (desc->params.nb_prev_steps != 5) { g_create_error = "only nb_prev_steps == 5 is implemented"; return RT_ERR_UNSUPPORTED; }
    if (desc->scheme == RT_SCHEME_LINE_OVERLAP && (!desc->overlap1 || !desc->overlap2)) { g_create_error = "overlap tables missing"; return RT_ERR_INVALID; }
    if ((desc->species_rule == RT_RULE_CH3OH || desc->species_rule == RT_RULE_OH_HF) && desc->nb_neutral_tables < 3) { g_create_error = "rule needs 3 neutral tables"; return RT_ERR_INVALID; }
    if (desc->species_rule == RT_RULE_H2CO && desc->nb_neutral_tables < 3) { g_create_error = "rule needs 3 neutral tables"; return RT_ERR_INVALID; }
    if (desc->species_rule == RT_RULE_NH3 && desc->nb_neutral_tables < 4) { g_create_error = "NH3 rule needs 4 neutral tables"; return RT_ERR_INVALID; }
    if (desc->species_rule == RT_RULE_H2O && desc->nb_neutral_tables < 6) { g_create_error = "H2O rule needs 6 neutral tables"; return RT_ERR_INVALID; }

    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { g_create_error = "no CUDA device: librt_b200 has no CPU fallback"; return RT_ERR_CUDA; }
    if (device < 0 || device >= ndev) { g_create_error = "device ordinal out of range"; return RT_ERR_INVALID; }

    rt_model* m = new rt_model;
    m->device = device;
    if (cudaSetDevice(device) != cudaSuccess) { g_create_error = "cudaSetDevice failed"; delete m; return RT_ERR_CUDA; }
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device);
    if (prop.major < 10) { g_create_error = "librt_b200 is built for sm_100a only"; delete m; return RT_ERR_CUDA; }
    // the critical chain of the LU (panels and what they wait for) runs on the model's stream and on the group streams, at the
    // highest stream priority; the far-column bulk work of the look-ahead schedule runs on side streams at the lowest
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
    if (getenv("RT_B200_NO_STREAM_PRIO")) prio_hi = prio_lo = 0;
    m->prio_hi = prio_hi; m->prio_lo = prio_lo;
    if (cudaStreamCreateWithPriority(&m->stream, cudaStreamNonBlocking, prio_hi) != cudaSuccess) { g_create_error = "stream"; delete m; return RT_ERR_CUDA; }
    if (cudaMallocHost((void**)&m->h_n_active, 4 * sizeof(int)) != cudaSuccess) { g_create_error = "pinned alloc"; delete m; return RT_ERR_CUDA; }
    if (cudaError_t ce = fused_kernels_init()) { g_create_error = std::string("cudaFuncSetAttribute failed: ") + cudaGetErrorString(ce); delete m; return RT_ERR_CUDA; }
    if (cudaError_t ce = lu_kernels_init()) { g_create_error = std::string("cudaFuncSetAttribute failed: ") + cudaGetErrorString(ce); delete m; return RT_ERR_CUDA; }

    m->desc = *desc;
    m->use_fused = getenv("RT_B200_NO_FUSED") == nullptr; // development switch: force the multi-kernel path
    m->mol_name = desc->mol_name ? desc->mol_name : "";
    DevModel& d = m->d;
    d.N = N;
    d.Np = (N + RT_NB - 1) / RT_NB * RT_NB;
    d.ld = d.Np + RT_NB;
    d.K = K; d.nb1 = desc->nb_neutral_tables; d.nb2 = K;
    d.rule = desc->species_rule; d.scheme = desc->scheme; d.ncomp = desc->dust_ncomp;
    d.prm = desc->params; d.mol_mass = desc->mol_mass;

    std::vector<int> vq(N, 0);
    std::vector<double> jq(N, 0.);
    if (desc->v) vq.assign(desc->v, desc->v + N);
    if (desc->j) jq.assign(desc->j, desc->j + N);
    d.energy = dev_upload(m, desc->energy, N, false);
    d.g = dev_upload(m, desc->g, N, false);
    d.v = dev_upload(m, vq.data(), N, false);
    d.jq = dev_upload(m, jq.data(), N, false);
    d.einst = dev_upload(m, desc->einst, (size_t)N * N, false);
    m->h_energy.assign(desc->energy, desc->energy + N);
    m->h_g.assign(desc->g, desc->g + N);
    m->h_einst.assign(desc->einst, desc->einst + (size_t)N * N);

    // ---- lines and groups ---------------------------------------------------------------------------------
    std::vector<int> tr_u, tr_l, grp;
    std::vector<double> dxn;
    auto add_line = [&](int u, int l) { tr_u.push_back(u); tr_l.push_back(l); return (int)tr_u.size() - 1; };
    if (desc->scheme == RT_SCHEME_LINE_OVERLAP) {
        std::vector<int> g5;
        if (desc->nb_line_groups > 0) g5.assign(desc->line_groups, desc->line_groups + 5 * (size_t)desc->nb_line_groups);
        else g5 = build_line_groups(desc);
        m->h_grp = g5;
        for (size_t i = 0; i < g5.size() / 5; i++) {
            const int* q = &g5[5 * i];
            const int la = add_line(q[1], q[2]);
            int lb = -1;
            double num = 0.;
            if (q[0] == 2) {
                lb = add_line(q[3], q[4]);
                // iteration_lvg.cpp:455, evaluated in the reference's order
                num = (desc->energy[q[1]] - desc->energy[q[2]] - desc->energy[q[3]] + desc->energy[q[4]]) * desc->params.speed_of_light;
            }
            grp.insert(grp.end(), {q[0], la, lb});
            dxn.push_back(num);
        }
    }
    else {
        // iteration_lvg.cpp:118-134: every pair i > j with A[i][j] > 0, in the reference's loop order
        for (int i = 1; i < N; i++)
            for (int j = 0; j < i; j++)
                if (desc->einst[(size_t)i * N + j] > 0.) {
                    const int la = add_line(i, j);
                    grp.insert(grp.end(), {1, la, -1});
                    dxn.push_back(0.);
                }
    }
    const int R = (int)tr_u.size();
    d.R = R; d.G = (int)grp.size() / 3;
    std::vector<double> Aul(R), Alu(R), E(R), sig((size_t)R * d.ncomp);
    for (int r = 0; r < R; r++) {
        const int u = tr_u[r], l = tr_l[r];
        if (u <= l || u >= N || l < 0) { g_create_error = "bad line in line_groups"; rt_model_destroy(m); return RT_ERR_INVALID; }
        Aul[r] = desc->einst[(size_t)u * N + l];
        Alu[r] = desc->einst[(size_t)l * N + u];
        E[r] = desc->energy[u] - desc->energy[l]; // iteration_lvg.cpp:167
        for (int c = 0; c < d.ncomp; c++) sig[(size_t)r * d.ncomp + c] = host_dust_sigma(desc->dust[c], E[r]);
    }
    m->h_tr_u = tr_u; m->h_tr_l = tr_l;
    for (int c = 0; c < d.ncomp; c++) {
        const rt_dust_comp& dc = desc->dust[c];
        m->h_dust_en.emplace_back(dc.ph_en, dc.ph_en + dc.nb_ph_en);
        m->h_dust_abs.emplace_back(dc.abs_coeff, dc.abs_coeff + dc.nb_ph_en);
        m->h_dust_exp.push_back(dc.wvl_exp);
    }
    d.tr_u = dev_upload(m, tr_u.data(), R, false); d.tr_l = dev_upload(m, tr_l.data(), R, false);
    d.tr_Aul = dev_upload(m, Aul.data(), R, false); d.tr_Alu = dev_upload(m, Alu.data(), R, false);
    d.tr_E = dev_upload(m, E.data(), R, false); d.tr_sigma = dev_upload(m, sig.data(), sig.size(), false);
    d.grp = dev_upload(m, grp.data(), grp.size(), false);
    d.grp_dxn = dev_upload(m, dxn.data(), dxn.size(), false);

    // per-row sparse lists (row CSR): (col, line << 1 | kind)
    std::vector<int> row_ptr(N + 1, 0);
    for (int r = 0; r < R; r++) { row_ptr[tr_u[r] + 1]++; row_ptr[tr_l[r] + 1]++; }
    for (int i = 0; i < N; i++) row_ptr[i + 1] += row_ptr[i];
    std::vector<int> fill(row_ptr.begin(), row_ptr.end() - 1), row_col(2 * (size_t)R), row_line(2 * (size_t)R);
    for (int r = 0; r < R; r++) {
        int e = fill[tr_u[r]]++; row_col[e] = tr_l[r]; row_line[e] = (r << 1);
        e = fill[tr_l[r]]++;     row_col[e] = tr_u[r]; row_line[e] = (r << 1) | 1;
    }
    d.row_ptr = dev_upload(m, row_ptr.data(), N + 1, false);
    d.row_col = dev_upload(m, row_col.data(), row_col.size(), false);
    d.row_line = dev_upload(m, row_line.data(), row_line.size(), false);

    // ---- collision tables, transposed to [jmax][imax] + linear derivatives (coll_rates.cpp:61-69) -----------
    for (int k = 0; k < K; k++) {
        const rt_coll_table& t = desc->tables[k];
        DevTable& T = d.tab[k];
        T.nb_lev = t.nb_lev; T.jmax = t.jmax; T.imax = t.nb_lev * (t.nb_lev - 1) / 2;
        if (t.jmax < 2 || t.nb_lev < 2 || t.nb_lev > N) { g_create_error = "bad collision table"; rt_model_destroy(m); return RT_ERR_INVALID; }
        T.max_temp = t.tgrid[t.jmax - 1];
        T.interp = t.interp;
        if (t.interp != RT_INTERP_LINEAR && t.interp != RT_INTERP_CUBIC_SPLINE) { g_create_error = "bad rt_coll_table::interp"; rt_model_destroy(m); return RT_ERR_INVALID; }
        std::vector<double> ct((size_t)T.imax * T.jmax), dt((size_t)T.imax * T.jmax, 0.);
        std::vector<double> u(T.jmax), d2(T.jmax);
        for (int i = 0; i < T.imax; i++) {
            const double* c = t.coeff + (size_t)i * T.jmax;
            const double* tg = t.tgrid;
            for (int j = 0; j < T.jmax; j++) {
                ct[(size_t)j * T.imax + i] = c[j];
                if (t.interp == RT_INTERP_LINEAR && j < T.jmax - 1)
                    dt[(size_t)j * T.imax + i] = (c[j + 1] - c[j]) / (tg[j + 1] - tg[j]);
            }
            if (t.interp == RT_INTERP_CUBIC_SPLINE) {
                // collision_data_cub_spline::calc_coeff_deriv, coll_rates.cpp:89-119: natural cubic spline, tridiagonal sweep
                // in the reference's expression order
                const int jm = T.jmax;
                d2[0] = u[0] = 0.;
                for (int j = 1; j < jm - 1; j++) {
                    const double sig = (tg[j] - tg[j - 1]) / (tg[j + 1] - tg[j - 1]);
                    const double p = sig * d2[j - 1] + 2.;
                    d2[j] = (sig - 1.) / p;
                    u[j] = (c[j + 1] - c[j]) / (tg[j + 1] - tg[j]) - (c[j] - c[j - 1]) / (tg[j] - tg[j - 1]);
                    u[j] = (6. * u[j] / (tg[j + 1] - tg[j - 1]) - sig * u[j - 1]) / p;
                }
                d2[jm - 1] = 0.;
                for (int j = jm - 2; j >= 0; j--) d2[j] = d2[j] * d2[j + 1] + u[j];
                for (int j = 0; j < jm; j++) dt[(size_t)j * T.imax + i] = d2[j];
            }
        }
        T.tgrid = dev_upload(m, t.tgrid, T.jmax, false);
        T.coeff_t = dev_upload(m, ct.data(), ct.size(), false);
        T.deriv_t = dev_upload(m, dt.data(), dt.size(), false);
    }

    // ---- escape-probability tables -------------------------------------------------------------------------
    d.lvg.nb_d = desc->lvg.nb_d; d.lvg.nb_g = desc->lvg.nb_g;
    d.lvg.delta = dev_upload(m, desc->lvg.delta, desc->lvg.nb_d, false);
    d.lvg.gamma = dev_upload(m, desc->lvg.gamma, desc->lvg.nb_g, false);
    d.lvg.p = dev_upload(m, desc->lvg.p, (size_t)desc->lvg.nb_d * desc->lvg.nb_g, false);
    if (desc->scheme == RT_SCHEME_LINE_OVERLAP) {
        const rt_overlap_table* src[2] = {desc->overlap1, desc->overlap2};
        for (int q = 0; q < 2; q++) {
            const rt_overlap_table& o = *src[q];
            std::vector<double> ld(o.nb_d);
            for (int i = 0; i < o.nb_d; i++) ld[i] = o.delta_is_log10 ? o.delta[i] : log10(o.delta[i]); // lvg_method_functions.cpp:293
            d.ov[q].nb_d = o.nb_d; d.ov[q].nb_dx = o.nb_dx; d.ov[q].nb_gr = o.nb_gr; d.ov[q].nb_g = o.nb_g;
            d.ov[q].log_delta = dev_upload(m, ld.data(), o.nb_d, false);
            d.ov[q].dx = dev_upload(m, o.dx, o.nb_dx, false);
            d.ov[q].gratio = dev_upload(m, o.gratio, o.nb_gr, false);
            d.ov[q].gamma = dev_upload(m, o.gamma, o.nb_g, false);
            d.ov[q].p = dev_upload(m, o.p, (size_t)o.nb_d * o.nb_dx * o.nb_gr * o.nb_g, false);
        }
        const rt_overlap_table &o1 = *src[0], &o2 = *src[1];
        d.ov_same_axes = o1.nb_d == o2.nb_d && o1.nb_dx == o2.nb_dx && o1.nb_gr == o2.nb_gr && o1.nb_g == o2.nb_g
                         && o1.delta_is_log10 == o2.delta_is_log10 && !memcmp(o1.delta, o2.delta, sizeof(double) * o1.nb_d)
                         && !memcmp(o1.dx, o2.dx, sizeof(double) * o1.nb_dx) && !memcmp(o1.gratio, o2.gratio, sizeof(double) * o1.nb_gr)
                         && !memcmp(o1.gamma, o2.gamma, sizeof(double) * o1.nb_g);
    }
    if (cudaDeviceSynchronize() != cudaSuccess || !m->last_error.empty()) {
        g_create_error = m->last_error.empty() ? "CUDA failure in rt_model_create" : m->last_error;
        rt_model_destroy(m);
        return RT_ERR_CUDA;
    }
    // the descriptor's pointers are not kept
    m->desc.energy = nullptr; m->desc.g = nullptr; m->desc.v = nullptr; m->desc.j = nullptr; m->desc.einst = nullptr;
    m->desc.tables = nullptr; m->desc.line_groups = nullptr; m->desc.overlap1 = m->desc.overlap2 = nullptr;
    m->desc.dust = nullptr; m->desc.mol_name = nullptr; m->desc.lvg.delta = m->desc.lvg.gamma = m->desc.lvg.p = nullptr;
    *out = m;
    return RT_OK;
}

static void free_cloud(rt_model* m)
{
    for (void* p : m->cloud_allocs) cudaFree(p);
    m->cloud_allocs.clear();
    m->has_cloud = false;
}

extern "C" rt_status rt_model_destroy(rt_model* m)
{
    if (!m) return RT_OK;
    cudaSetDevice(m->device);
    if (m->stream) cudaStreamSynchronize(m->stream);
    free_cloud(m);
    for (void* p : m->allocs) cudaFree(p);
    for (auto& e : m->ev_pool) cudaEventDestroy(e);
    for (auto& e : m->timer_ev) if (e) cudaEventDestroy(e);
    if (m->h_n_active) cudaFreeHost(m->h_n_active);
    for (int g = 0; g < RT_LU_MAX_GROUPS; g++) {
        if (m->lu_side[g]) cudaStreamDestroy(m->lu_side[g]);
        for (int e = 0; e < 3; e++) if (m->lu_ev[e][g]) cudaEventDestroy(m->lu_ev[e][g]);
    }
    for (int g = 1; g < RT_LU_MAX_GROUPS; g++) {
        if (m->lu_streams[g]) cudaStreamDestroy(m->lu_streams[g]);
        if (m->lu_join[g]) cudaEventDestroy(m->lu_join[g]);
    }
    if (m->lu_fork) cudaEventDestroy(m->lu_fork);
    if (m->stream) cudaStreamDestroy(m->stream);
    delete m;
    return RT_OK;
}

// ---- profiling hooks ------------------------------------------------------------------------------------------
int rt_prof_class(rt_model* m, const char* name)
{
    for (size_t i = 0; i < m->prof_names.size(); i++)
        if (m->prof_names[i] == name) return (int)i;
    m->prof_names.push_back(name);
    m->prof_ms.push_back(0.);
    m->prof_launches.push_back(0);
    m->prof_flops.push_back(0.);
    m->prof_bytes.push_back(0.);
    return (int)m->prof_names.size() - 1;
}
static cudaEvent_t get_event(rt_model* m)
{
    if (!m->ev_pool.empty()) { cudaEvent_t e = m->ev_pool.back(); m->ev_pool.pop_back(); return e; }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
}
void rt_prof_begin(rt_model* m, int cls, cudaStream_t s, double flops_per_layer, double bytes_per_layer)
{
    m->prof_launches[cls]++;
    m->prof_flops[cls] += flops_per_layer * m->prof_layers;
    m->prof_bytes[cls] += bytes_per_layer * m->prof_layers;
    if (!m->prof_timing) return;
    cudaEvent_t a = get_event(m), b = get_event(m);
    cudaEventRecord(a, s);
    m->prof_pending.push_back({cls, a, b});
}
void rt_prof_end(rt_model* m, int cls, cudaStream_t s)
{
    if (!m->prof_timing) return;
    cudaEventRecord(m->prof_pending.back().b, s);
}
static void prof_collect(rt_model* m)
{
    for (auto& p : m->prof_pending) {
        float ms = 0.f;
        cudaEventSynchronize(p.b);
        cudaEventElapsedTime(&ms, p.a, p.b);
        m->prof_ms[p.cls] += ms;
        m->ev_pool.push_back(p.a);
        m->ev_pool.push_back(p.b);
    }
    m->prof_pending.clear();
}

// collision_data::locate, coll_rates.cpp:71-82
static int coll_locate(const double* tgrid, int jmax, double temp)
{
    int l = 0, r = jmax - 1;
    while (r - l > 1) {
        const int j = l + ((r - l) >> 1);
        if (tgrid[j] < temp) l = j; else r = j;
    }
    return l;
}

extern "C" rt_status rt_cloud_upload(rt_model* m, const rt_layer_soa* lay)
{
    if (!m || !lay || lay->nb_lay < 1) return RT_ERR_INVALID;
    m->last_error.clear(); // errors are per call: a recoverable error of an earlier call must not fail this one
    if (lay->nb_lay > 65535) { m->last_error = "more than 65535 layers in one cloud (layers ride in gridDim.y): split the stack"; return RT_ERR_UNSUPPORTED; }
    if (lay->dust_ncomp != m->d.ncomp) { m->last_error = "inconsistency in the interstellar dust data"; return RT_ERR_DUST_INCONSISTENT; }
    std::vector<int> chain_first = {0, lay->nb_lay};
    if (lay->nb_models > 0) {
        bool ok = lay->model_first && lay->model_first[0] == 0 && lay->model_first[lay->nb_models] == lay->nb_lay;
        for (int k = 0; ok && k < lay->nb_models; k++) ok = lay->model_first[k + 1] > lay->model_first[k];
        if (!ok) { m->last_error = "rt_layer_soa::model_first must ascend from 0 to nb_lay"; return RT_ERR_INVALID; }
        chain_first.assign(lay->model_first, lay->model_first + lay->nb_models + 1);
    }
    CK(cudaSetDevice(m->device));
    CK(cudaStreamSynchronize(m->stream));
    DevModel& d = m->d;
    const int L = lay->nb_lay, N = d.N, Np = d.Np, K = d.K;
    m->chain_first = chain_first;
    m->cloud_reuse = m->has_cloud && d.L == L;
    m->cloud_slot = 0;
    if (!m->cloud_reuse) free_cloud(m);
    m->has_cloud = false; // until this upload completes
    d.L = L;
    // the three per-layer setters, radiative_transfer.cpp:238-242
    std::vector<double> vw(L), conc((size_t)L * K, 0.);
    std::vector<int> tidx((size_t)L * K, 0);
    // temperature grids are needed on the host for locate(); fetch them back from the tables
    std::vector<std::vector<double>>& tg = m->h_tgrid;
    if ((int)tg.size() != K) {
        tg.assign(K, {});
        for (int k = 0; k < K; k++) {
            tg[k].resize(d.tab[k].jmax);
            CK(cudaMemcpy(tg[k].data(), d.tab[k].tgrid, sizeof(double) * d.tab[k].jmax, cudaMemcpyDeviceToHost));
        }
    }
    for (int l = 0; l < L; l++) {
        // iteration_lvg.cpp:65
        vw[l] = lay->vel_width ? lay->vel_width[l]
                               : pow(2. * d.prm.boltzmann_constant * lay->temp_n[l] / d.mol_mass + lay->vel_turb[l] * lay->vel_turb[l], 0.5);
        double* c = &conc[(size_t)l * K];
        int* ix = &tidx[(size_t)l * K];
        // set_gas_param: coll_rates.cpp:152-174 + coll_rates_ch3oh.cpp:472-481, coll_rates_oh.cpp:380-389, coll_rates_h2o.cpp:515-527
        for (int k = 0; k < d.nb1; k++) ix[k] = coll_locate(tg[k].data(), d.tab[k].jmax, lay->temp_n[l]);
        for (int k = d.nb1; k < d.nb2; k++) { c[k] = lay->el_conc ? lay->el_conc[l] : 0.; ix[k] = coll_locate(tg[k].data(), d.tab[k].jmax, lay->temp_el[l]); }
        if (lay->partner_conc) {
            for (int k = 0; k < K; k++) c[k] = lay->partner_conc[(size_t)l * K + k];
            continue;
        }
        const double he = lay->he_conc[l], ph2 = lay->ph2_conc[l], oh2 = lay->oh2_conc[l], h = lay->h_conc[l];
        switch (d.rule) {
        case RT_RULE_CH3OH: case RT_RULE_OH_HF: case RT_RULE_H2CO: c[0] = he; c[1] = ph2; c[2] = oh2; break;
        case RT_RULE_NH3: c[0] = he; c[1] = ph2; c[2] = oh2; c[3] = h; break; // coll_rates_nh3.cpp:261-271
        case RT_RULE_H2O: c[0] = he; c[1] = he + 0.2 * h; c[2] = ph2; c[3] = oh2; c[4] = ph2 + oh2; c[5] = h; break;
        default: { const double cc[4] = {he, ph2, oh2, h}; for (int k = 0; k < d.nb1; k++) c[k] = cc[k < 4 ? k : 3]; }
        }
    }
    d.temp_n = dev_upload(m, lay->temp_n, L, true);
    d.temp_el = dev_upload(m, lay->temp_el, L, true);
    d.mol_conc = dev_upload(m, lay->mol_conc, L, true);
    d.vel_grad = dev_upload(m, lay->velg_n, L, true);
    d.vel_width = dev_upload(m, vw.data(), L, true);
    d.dust_conc = dev_upload(m, lay->dust_conc, (size_t)L * d.ncomp, true);
    d.conc = dev_upload(m, conc.data(), conc.size(), true);
    d.tidx = dev_upload(m, tidx.data(), tidx.size(), true);
    m->h_dz.assign(lay->dz ? lay->dz : lay->temp_n, (lay->dz ? lay->dz : lay->temp_n) + L);
    m->h_vel_n.assign(lay->vel_n ? lay->vel_n : lay->temp_n, (lay->vel_n ? lay->vel_n : lay->temp_n) + L);
    m->h_height = lay->height;
    m->h_vel_turb.assign(lay->vel_turb, lay->vel_turb + L);
    m->h_h2_conc.resize(L);
    for (int l = 0; l < L; l++) m->h_h2_conc[l] = lay->ph2_conc[l] + lay->oh2_conc[l]; // maser_luminosity.cpp:79,95
    m->h_mol_conc.assign(lay->mol_conc, lay->mol_conc + L);
    m->h_velg.assign(lay->velg_n, lay->velg_n + L);

    const size_t mat = (size_t)L * Np * Np;
    d.Cmat = dev_alloc<double>(m, mat, true, false);
    d.Mmat = dev_alloc<double>(m, (size_t)L * Np * d.ld, true, false);
    d.pop = dev_alloc<double>(m, (size_t)L * N, true);
    d.rhs = dev_alloc<double>(m, (size_t)L * Np, true);
    d.hist_pop = dev_alloc<double>(m, (size_t)L * RT_HIST * N, true);
    d.hist_res = dev_alloc<double>(m, (size_t)L * RT_HIST * N, true);
    d.opt = dev_alloc<double>(m, (size_t)L * N, true);
    d.yd = dev_alloc<double>(m, (size_t)L * std::max(1, d.R), true);
    d.yu = dev_alloc<double>(m, (size_t)L * std::max(1, d.R), true);
    d.ipiv = dev_alloc<int>(m, (size_t)L * Np, true);
    d.lu_aff = dev_alloc<int>(m, (size_t)L * (Np / RT_NB) * RT_AFF_STRIDE, true);
    int** ip[] = {&d.iter_nb, &d.nb_after, &d.head_pop, &d.head_res, &d.n_hist_pop, &d.n_hist_res, &d.conv, &d.active_flag, &d.active};
    for (int** p : ip) *p = dev_alloc<int>(m, L, true);
    double** dp[] = {&d.best_eq, &d.eq_err, &d.pop_err, &d.rel_err};
    for (double** p : dp) *p = dev_alloc<double>(m, L, true);
    d.eq_bits = dev_alloc<unsigned long long>(m, L, true);
    d.n_active = dev_alloc<int>(m, 4, true);
    d.lay_list = dev_alloc<int>(m, L, true);
    if (!d.Cmat || !d.Mmat || !d.pop || !d.hist_pop || !d.hist_res || !d.n_active || !m->last_error.empty()) {
        m->cloud_reuse = false;
        free_cloud(m);
        return RT_ERR_NOMEM;
    }
    m->cloud_reuse = false;
    m->h_status.assign(L, rt_layer_status{});
    // K0: collisional matrix of every layer, once (it does not depend on the populations)
    const int c0 = rt_prof_class(m, "coll_matrix");
    m->has_cloud = true;
    return rt_cloud_rebuild(m);
}

// K0 from the layer arrays already resident in HBM: collisional matrix of every layer, once per cloud
// (it does not depend on the populations)
extern "C" rt_status rt_cloud_rebuild(rt_model* m)
{
    if (!m) return RT_ERR_INVALID;
    m->last_error.clear();
    if (!m->has_cloud) { m->last_error = "no cloud uploaded"; return RT_ERR_NO_CLOUD; }
    CK(cudaSetDevice(m->device));
    const DevModel& d = m->d;
    const int c0 = rt_prof_class(m, "coll_matrix");
    m->prof_layers = d.L;
    double tab_bytes = 0.;
    for (int k = 0; k < d.K; k++) tab_bytes += 16. * d.tab[k].imax; // coeff + deriv at the layer's temperature index
    rt_prof_begin(m, c0, m->stream, 0., tab_bytes + 8. * d.N * d.N);
    launch_coll_matrix(d, m->stream, d.Cmat, 0);
    rt_prof_end(m, c0, m->stream);
    CK(cudaStreamSynchronize(m->stream));
    CK(cudaGetLastError());
    return RT_OK;
}

// ---- internal building blocks ---------------------------------------------------------------------------------
// one calc_new_pop for the active list: J, M, df  ->  solve M x = e0  (iteration_lvg.cpp:87-110)
static void enqueue_calc_new_pop(rt_model* m, int bound, double* df_out)
{
    const DevModel& d = m->d;
    cudaStream_t s = m->stream;
    const int c_int = rt_prof_class(m, "intensity"), c_asm = rt_prof_class(m, "assemble");
    // algorithmic bytes per layer: line records {u,l: int32; A_ul,A_lu,E,sigma: f64} + 2 populations in, yd/yu out
    rt_prof_begin(m, c_int, s, 0., (40. + 16. + 16.) * d.R);
    launch_intensity(d, s, bound);
    rt_prof_end(m, c_int, s);
    // read C [N][N] + write M [N][N] + pop in + df/rhs out + the sparse radiative list (yd,yu: 16 R; row CSR: 16 R)
    rt_prof_begin(m, c_asm, s, 2. * d.N * d.N + 4. * d.R, 16. * d.N * d.N + 16. * d.N + 32. * d.R);
    launch_assemble(d, s, bound, df_out);
    rt_prof_end(m, c_asm, s);
    launch_lu_solve(m, s, bound);
}

// start vectors for the flagged + compacted layers
static void enqueue_start(rt_model* m, int start_mode, int bound)
{
    const DevModel& d = m->d;
    cudaStream_t s = m->stream;
    if (start_mode == RT_START_BOUNDARY) {
        // boundary_layer_populations, iteration_control.cpp:52-91: neutrals + 0.5 A, row 0 = 1, solve
        const int c0 = rt_prof_class(m, "boundary_matrix");
        double tab_bytes = 0.;
        for (int k = 0; k < d.nb1; k++) tab_bytes += 16. * d.tab[k].imax;
        rt_prof_begin(m, c0, s, 0., tab_bytes + 12. * d.N * d.N);
        launch_coll_matrix(d, s, d.Mmat, 1, bound); // only the layers this start is for
        rt_prof_end(m, c0, s);
        launch_set_row0(d, s, bound);
        launch_set_rhs(d, s, bound);
        launch_lu_solve(m, s, bound);
        launch_copy_rhs_to_pop(d, s, bound);
    }
    else if (start_mode == RT_START_LTE) launch_boltzmann(d, s);
}

// iteration_control::calculate_populations for the layers whose active_flag is set (the caller flagged them: a range,
// the failed layers of a range, or an explicit list); `n_flagged_hint` layers at most
static rt_status solve_flagged(rt_model* m, int start_mode, int max_iter, double min_err, int accel)
{
    const DevModel& d = m->d;
    cudaStream_t s = m->stream;
    launch_compact(d, s);
    CK(cudaMemcpyAsync(m->h_n_active, d.n_active, sizeof(int), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    int bound = m->h_n_active[0];
    if (bound == 0) return RT_OK;
    const int n_layers = bound;
    m->prof_layers = bound;
    enqueue_start(m, start_mode, bound);
    launch_ctl_init_flagged(d, s, accel);
    if (m->use_fused && fused_supported(d)) {
        // small molecules: the whole iteration of a layer runs inside one persistent CTA (fused_kernels.cu)
        const int c_f = rt_prof_class(m, "fused_solve");
        rt_prof_begin(m, c_f, s, 0., 0.);
        launch_fused_solve(d, s, bound, accel, max_iter, min_err);
        rt_prof_end(m, c_f, s);
        CK(cudaStreamSynchronize(s));
        CK(cudaGetLastError());
        // the active list is untouched by the fused kernel: it still names the layers of this solve
        std::vector<int> itn(d.L), act(n_layers);
        CK(cudaMemcpy(itn.data(), d.iter_nb, sizeof(int) * d.L, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(act.data(), d.active, sizeof(int) * n_layers, cudaMemcpyDeviceToHost));
        int mx = 0;
        double layer_iters = 0.;
        for (int l : act) { mx = std::max(mx, itn[l]); layer_iters += itn[l]; }
        m->iters_enqueued += mx;
        // algorithmic work of the launch, known only now: per layer-iteration one LU + two triangular solves +
        // the residual product; HBM bytes only when C(layer) does not fit beside M in shared memory (Np > 96)
        const double n = d.N;
        m->prof_flops[c_f] += layer_iters * (2. / 3. * n * n * n + 4. * n * n);
        m->prof_bytes[c_f] += layer_iters * (d.Np > 96 ? 8. * n * n : 0.) + (double)n_layers * 8. * n * n;
        return RT_OK;
    }
    const int c_ctl = rt_prof_class(m, "controller");
    // n_active after iteration `it` lands in the pinned ring slot (it % RT_RING) + 1; the host reads it one
    // iteration late, so the GPU always has the next iteration queued while the host looks at the count
    cudaEvent_t ev[2];
    CK(cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming));
    int it = 0;
    for (; it < max_iter && bound > 0; it++) {
        // rt_internal.h (rt_launch): on while the batch is small enough to be latency-bound -- measured in layers of N = 768
        g_rt_pdl_on = !m->prof_timing && (double)bound * d.Np * d.Np <= (double)g_rt_pdl_max_layers * 768. * 768.;
        enqueue_calc_new_pop(m, bound, nullptr);
        rt_prof_begin(m, c_ctl, s, 0., 8. * 6. * d.N);
        launch_ctl_step(d, s, bound, accel, max_iter, min_err);
        launch_compact(d, s);
        rt_prof_end(m, c_ctl, s);
        CK(cudaMemcpyAsync(m->h_n_active + 1 + (it & 1), d.n_active, sizeof(int), cudaMemcpyDeviceToHost, s));
        CK(cudaEventRecord(ev[it & 1], s));
        if (it >= 1) {
            CK(cudaEventSynchronize(ev[(it - 1) & 1]));
            bound = m->h_n_active[1 + ((it - 1) & 1)]; // n_active only shrinks: a stale value is a valid bound
            m->prof_layers = bound;
        }
    }
    g_rt_pdl_on = 0;
    CK(cudaStreamSynchronize(s));
    cudaEventDestroy(ev[0]);
    cudaEventDestroy(ev[1]);
    CK(cudaGetLastError());
    m->iters_enqueued += it;
    return RT_OK;
}

// ... for layers [lb, le) (or only the non-converged ones of that range)
static rt_status solve_range(rt_model* m, int lb, int le, int start_mode, int max_iter, double min_err, int accel, bool only_failed)
{
    if (only_failed) launch_set_active_failed(m->d, m->stream, lb, le);
    else launch_set_active_range(m->d, m->stream, lb, le);
    return solve_flagged(m, start_mode, max_iter, min_err, accel);
}

// flags <- the listed layers (host list -> d.lay_list -> active_flag)
static rt_status flag_layers(rt_model* m, const std::vector<int>& lay)
{
    if (!lay.empty()) CK(cudaMemcpyAsync(m->d.lay_list, lay.data(), sizeof(int) * lay.size(), cudaMemcpyHostToDevice, m->stream));
    launch_set_active_list(m->d, m->stream, (int)lay.size());
    CK(cudaStreamSynchronize(m->stream)); // `lay` is pageable host memory that the caller may change next
    return RT_OK;
}

// One attempt of one chain step (radiative_transfer.cpp:247-257 / :263-272): the listed layers get the populations of the
// layer before them when that one converged (`warm`), boundary_layer_populations otherwise (`cold`), then iterate together.
static rt_status chain_attempt(rt_model* m, const std::vector<int>& cold, const std::vector<int>& warm, int max_iter, double min_err, int accel)
{
    const DevModel& d = m->d;
    rt_status r;
    if (!cold.empty()) {
        if ((r = flag_layers(m, cold))) return r;
        launch_compact(d, m->stream);
        m->prof_layers = (int)cold.size();
        enqueue_start(m, RT_START_BOUNDARY, (int)cold.size());
    }
    if (!warm.empty()) {
        CK(cudaMemcpyAsync(d.lay_list, warm.data(), sizeof(int) * warm.size(), cudaMemcpyHostToDevice, m->stream));
        launch_copy_prev_pop(d, m->stream, (int)warm.size());
        CK(cudaStreamSynchronize(m->stream));
    }
    std::vector<int> all(cold);
    all.insert(all.end(), warm.begin(), warm.end());
    if ((r = flag_layers(m, all))) return r;
    return solve_flagged(m, RT_START_GIVEN, max_iter, min_err, accel);
}

static bool is_methanol(const rt_model* m) { return m->mol_name == "CH3OHa" || m->mol_name == "CH3OHe"; }

static rt_status fetch_status(rt_model* m, rt_layer_status* st)
{
    if (!st) return RT_OK;
    const DevModel& d = m->d;
    const int L = d.L;
    std::vector<int> it(L), cv(L);
    std::vector<double> e1(L), e2(L), e3(L);
    CK(cudaMemcpy(it.data(), d.iter_nb, sizeof(int) * L, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(cv.data(), d.conv, sizeof(int) * L, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(e1.data(), d.eq_err, sizeof(double) * L, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(e2.data(), d.pop_err, sizeof(double) * L, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(e3.data(), d.rel_err, sizeof(double) * L, cudaMemcpyDeviceToHost));
    for (int l = 0; l < L; l++) st[l] = rt_layer_status{it[l], cv[l], e1[l], e2[l], e3[l]};
    return RT_OK;
}

#define NEED_CLOUD()                                                   \
    if (!m) return RT_ERR_INVALID;                                     \
    m->last_error.clear();                                             \
    if (!m->has_cloud) { m->last_error = "no cloud uploaded"; return RT_ERR_NO_CLOUD; } \
    CK(cudaSetDevice(m->device))

extern "C" {

rt_status rt_set_populations(rt_model* m, const double* pop)
{
    NEED_CLOUD();
    CK(cudaMemcpyAsync(m->d.pop, pop, sizeof(double) * (size_t)m->d.L * m->d.N, cudaMemcpyHostToDevice, m->stream));
    CK(cudaStreamSynchronize(m->stream));
    return RT_OK;
}

rt_status rt_get_populations(rt_model* m, double* pop)
{
    NEED_CLOUD();
    CK(cudaMemcpyAsync(pop, m->d.pop, sizeof(double) * (size_t)m->d.L * m->d.N, cudaMemcpyDeviceToHost, m->stream));
    CK(cudaStreamSynchronize(m->stream));
    return RT_OK;
}

rt_status rt_copy_populations_device(rt_model* m, double* dst_device)
{
    NEED_CLOUD();
    if (!dst_device) return RT_ERR_INVALID;
    CK(cudaMemcpyAsync(dst_device, m->d.pop, sizeof(double) * (size_t)m->d.L * m->d.N, cudaMemcpyDeviceToDevice, m->stream));
    CK(cudaStreamSynchronize(m->stream));
    return RT_OK;
}

rt_status rt_get_status(rt_model* m, rt_layer_status* st)
{
    NEED_CLOUD();
    return fetch_status(m, st);
}

rt_status rt_solve_device(rt_model* m, int start_mode, int max_iter, double min_err, int accel)
{
    NEED_CLOUD();
    if (start_mode == RT_START_CHAINED) { m->last_error = "use rt_calc_molecular_populations for the chained start"; return RT_ERR_INVALID; }
    return solve_range(m, 0, m->d.L, start_mode, max_iter, min_err, accel, false);
}

rt_status rt_solve(rt_model* m, double* pop, int start_mode, int max_iter, double min_err, int accel, rt_layer_status* st)
{
    NEED_CLOUD();
    if (!pop) return RT_ERR_INVALID;
    if (start_mode == RT_START_GIVEN) { rt_status r = rt_set_populations(m, pop); if (r) return r; }
    rt_status r = rt_solve_device(m, start_mode, max_iter, min_err, accel);
    if (r) return r;
    r = rt_get_populations(m, pop);
    if (r) return r;
    return fetch_status(m, st);
}

rt_status rt_calc_molecular_populations(rt_model* m, double* pop, int start_mode, int acceleration, int max_iter_acc,
                                        int max_iter_ext, double min_err, rt_layer_status* st)
{
    NEED_CLOUD();
    if (!pop) return RT_ERR_INVALID;
    const DevModel& d = m->d;
    const int L = d.L, N = d.N;
    const int iter_max = acceleration ? max_iter_acc : max_iter_ext; // radiative_transfer.cpp:229-231
    const bool retry = acceleration && !is_methanol(m);              // :259
    rt_status r;
    if (start_mode != RT_START_CHAINED) {
        if (start_mode == RT_START_GIVEN && (r = rt_set_populations(m, pop))) return r;
        if ((r = solve_range(m, 0, L, start_mode, iter_max, min_err, acceleration, false))) return r;
        if (retry) {
            // same kind of start for the failed layers (:263-269), un-accelerated, extended iteration cap
            if (start_mode == RT_START_GIVEN) {
                std::vector<int> cv(L);
                CK(cudaMemcpy(cv.data(), d.conv, sizeof(int) * L, cudaMemcpyDeviceToHost));
                for (int l = 0; l < L; l++)
                    if (!cv[l]) CK(cudaMemcpyAsync(d.pop + (size_t)l * N, pop + (size_t)l * N, sizeof(double) * N, cudaMemcpyHostToDevice, m->stream));
            }
            if ((r = solve_range(m, 0, L, start_mode, max_iter_ext, min_err, 0, true))) return r;
        }
    }
    else {
        // warm-start chain, radiative_transfer.cpp:236-281: layer l starts from layer l-1 if that one converged.  A stacked
        // cloud (rt_layer_soa::model_first) holds one chain per model; chain step l of ALL models is one batch.
        const std::vector<int>& first = m->chain_first;
        const int nm = (int)first.size() - 1;
        int maxlen = 0;
        for (int k = 0; k < nm; k++) maxlen = std::max(maxlen, first[k + 1] - first[k]);
        std::vector<int> prev_found(nm, 0), conv(L);
        for (int step = 0; step < maxlen; step++) {
            std::vector<int> cold, warm;
            for (int k = 0; k < nm; k++) {
                const int lay = first[k] + step;
                if (lay >= first[k + 1]) continue;
                (step > 0 && prev_found[k] ? warm : cold).push_back(lay);
            }
            if ((r = chain_attempt(m, cold, warm, iter_max, min_err, acceleration))) return r;
            CK(cudaMemcpy(conv.data(), d.conv, sizeof(int) * L, cudaMemcpyDeviceToHost));
            if (retry) { // the same kind of start once more, un-accelerated (:259-276)
                std::vector<int> cold2, warm2;
                for (int lay : cold) if (!conv[lay]) cold2.push_back(lay);
                for (int lay : warm) if (!conv[lay]) warm2.push_back(lay);
                if (!cold2.empty() || !warm2.empty()) {
                    if ((r = chain_attempt(m, cold2, warm2, max_iter_ext, min_err, 0))) return r;
                    CK(cudaMemcpy(conv.data(), d.conv, sizeof(int) * L, cudaMemcpyDeviceToHost));
                }
            }
            for (int k = 0; k < nm; k++) {
                const int lay = first[k] + step;
                if (lay < first[k + 1]) prev_found[k] = conv[lay];
            }
        }
    }
    if ((r = rt_get_populations(m, pop))) return r;
    return fetch_status(m, st);
}

rt_status rt_boundary_populations(rt_model* m, double* pop)
{
    NEED_CLOUD();
    const DevModel& d = m->d;
    launch_set_active_range(d, m->stream, 0, d.L);
    launch_compact(d, m->stream);
    enqueue_start(m, RT_START_BOUNDARY, d.L);
    return rt_get_populations(m, pop);
}

rt_status rt_assemble(rt_model* m, const double* pop, double* matrix, double* df)
{
    NEED_CLOUD();
    const DevModel& d = m->d;
    const int L = d.L, N = d.N, Np = d.Np;
    rt_status r = rt_set_populations(m, pop);
    if (r) return r;
    DevScratch scratch(m->stream);
    double* df_dev = (double*)scratch.get(sizeof(double) * (size_t)L * N);
    NEED_BUF(df_dev);
    launch_set_active_range(d, m->stream, 0, L);
    launch_compact(d, m->stream);
    CK(cudaMemsetAsync(d.eq_bits, 0, sizeof(unsigned long long) * L, m->stream));
    launch_intensity(d, m->stream, L);
    launch_assemble(d, m->stream, L, df_dev);
    CK(cudaStreamSynchronize(m->stream));
    if (df) CK(cudaMemcpy(df, df_dev, sizeof(double) * (size_t)L * N, cudaMemcpyDeviceToHost));
    if (matrix) // strip the padding, layer by layer
        for (int l = 0; l < L; l++)
            CK(cudaMemcpy2D(matrix + (size_t)l * N * N, sizeof(double) * N, d.Mmat + (size_t)l * Np * d.ld, sizeof(double) * d.ld,
                            sizeof(double) * N, N, cudaMemcpyDeviceToHost));
    return RT_OK;
}

rt_status rt_step(rt_model* m, const double* old_pop, double* new_pop, double* eq_err)
{
    NEED_CLOUD();
    const DevModel& d = m->d;
    const int L = d.L, N = d.N, Np = d.Np;
    rt_status r = rt_set_populations(m, old_pop);
    if (r) return r;
    launch_set_active_range(d, m->stream, 0, L);
    launch_compact(d, m->stream);
    CK(cudaMemsetAsync(d.eq_bits, 0, sizeof(unsigned long long) * L, m->stream));
    enqueue_calc_new_pop(m, L, nullptr);
    CK(cudaStreamSynchronize(m->stream));
    CK(cudaGetLastError());
    if (new_pop) CK(cudaMemcpy2D(new_pop, sizeof(double) * N, d.rhs, sizeof(double) * Np, sizeof(double) * N, L, cudaMemcpyDeviceToHost));
    if (eq_err) CK(cudaMemcpy(eq_err, d.eq_bits, sizeof(double) * L, cudaMemcpyDeviceToHost)); // bits of a non-negative double
    return RT_OK;
}

rt_status rt_profile_reset(rt_model* m, int enable_timing)
{
    if (!m) return RT_ERR_INVALID;
    prof_collect(m);
    std::fill(m->prof_ms.begin(), m->prof_ms.end(), 0.);
    std::fill(m->prof_launches.begin(), m->prof_launches.end(), 0);
    std::fill(m->prof_flops.begin(), m->prof_flops.end(), 0.);
    std::fill(m->prof_bytes.begin(), m->prof_bytes.end(), 0.);
    m->prof_timing = enable_timing != 0;
    m->iters_enqueued = 0;
    return RT_OK;
}

rt_status rt_profile_get(rt_model* m, int cap, const char** names, double* ms, long long* launches, int* n)
{
    if (!m || !n) return RT_ERR_INVALID;
    prof_collect(m);
    const int k = (int)std::min<size_t>(cap, m->prof_names.size());
    for (int i = 0; i < k; i++) {
        if (names) names[i] = m->prof_names[i].c_str();
        if (ms) ms[i] = m->prof_ms[i];
        if (launches) launches[i] = m->prof_launches[i];
    }
    *n = k;
    return RT_OK;
}

long long rt_iterations_enqueued(const rt_model* m) { return m ? m->iters_enqueued : 0; }
long long rt_launch_count(void) { return g_rt_launches.load(); }

rt_status rt_profile_get_work(rt_model* m, int cap, double* flops, double* bytes)
{
    if (!m) return RT_ERR_INVALID;
    const int k = (int)std::min<size_t>(cap, m->prof_names.size());
    for (int i = 0; i < k; i++) {
        if (flops) flops[i] = m->prof_flops[i];
        if (bytes) bytes[i] = m->prof_bytes[i];
    }
    return RT_OK;
}

rt_status rt_timer_begin(rt_model* m)
{
    if (!m) return RT_ERR_INVALID;
    CK(cudaSetDevice(m->device));
    for (int i = 0; i < 2; i++)
        if (!m->timer_ev[i]) CK(cudaEventCreate(&m->timer_ev[i]));
    CK(cudaStreamSynchronize(m->stream));
    CK(cudaEventRecord(m->timer_ev[0], m->stream));
    return RT_OK;
}

rt_status rt_timer_end(rt_model* m, double* ms)
{
    if (!m || !ms || !m->timer_ev[0]) return RT_ERR_INVALID;
    CK(cudaEventRecord(m->timer_ev[1], m->stream));
    CK(cudaEventSynchronize(m->timer_ev[1]));
    float f = 0.f;
    CK(cudaEventElapsedTime(&f, m->timer_ev[0], m->timer_ev[1]));
    *ms = f;
    return RT_OK;
}

} // extern "C"

// ---- rt_intensity / rt_maser_postpass ---------------------------------------------------------------------------
struct PostArgs {
    const double* pop;
    const int *up, *low;
    const double* sigma;
    const double *dz, *vel_n, *vgrid, *vel_turb;
    double height, delta_aspect, gw_offset;
    int nb_freq, nb_aspect, gw_up, gw_low;
    double *inv_arr, *gain_arr, *scalars, *tau_aspect, *tau_freq;
};
void launch_postpass_impl(const DevModel& d, cudaStream_t s, const PostArgs& a, int nb_trans);

static std::vector<double> line_sigma(const rt_model* m, double energy)
{
    std::vector<double> s(m->d.ncomp);
    for (int c = 0; c < m->d.ncomp; c++) {
        rt_dust_comp dc;
        dc.nb_ph_en = (int)m->h_dust_en[c].size(); dc.ph_en = m->h_dust_en[c].data(); dc.abs_coeff = m->h_dust_abs[c].data();
        dc.wvl_exp = m->h_dust_exp[c];
        s[c] = host_dust_sigma(dc, energy);
    }
    return s;
}

extern "C" rt_status rt_intensity(rt_model* m, int lay, int up, int low, const double* pop, double* intensity)
{
    NEED_CLOUD();
    const DevModel& d = m->d;
    const int N = d.N;
    if (lay < 0 || lay >= d.L || up < 0 || up >= N || low < 0 || low >= N || !pop || !intensity) return RT_ERR_INVALID;
    int which = 0, u1 = up, l1 = low, u2 = up, l2 = low;
    if (d.scheme == RT_SCHEME_LINE_OVERLAP) {
        // iteration_lvg.cpp:503-526: look the line up in line_list; not found -> 0
        bool found = false;
        for (size_t i = 0; i < m->h_grp.size() / 5 && !found; i++) {
            const int* g = &m->h_grp[5 * i];
            for (int k = 0; k < g[0]; k++)
                if (g[2 + 2 * k] == low && g[1 + 2 * k] == up) {
                    found = true;
                    if (g[0] == 2) { which = k + 1; u1 = g[1]; l1 = g[2]; u2 = g[3]; l2 = g[4]; }
                    break;
                }
        }
        if (!found) { *intensity = 0.; return RT_OK; }
    }
    const double e1 = m->h_energy[u1] - m->h_energy[l1];
    std::vector<double> buf(1 + RT_MAX_DUST, 0.);
    const std::vector<double> sg = line_sigma(m, e1);
    for (int c = 0; c < d.ncomp; c++) buf[1 + c] = sg[c];
    DevScratch scratch(m->stream);
    double* pop_dev = (double*)scratch.get(sizeof(double) * N, pop);
    double* out_dev = (double*)scratch.get(sizeof(double) * buf.size(), buf.data());
    NEED_BUF(pop_dev && out_dev);
    const std::vector<double>& A = m->h_einst;
    launch_intensity_single(d, m->stream, lay, u1, l1, A[(size_t)u1 * N + l1], A[(size_t)l1 * N + u1], u2, l2,
                            A[(size_t)u2 * N + l2], A[(size_t)l2 * N + u2], which, pop_dev, out_dev);
    CK(cudaStreamSynchronize(m->stream));
    CK(cudaMemcpy(intensity, out_dev, sizeof(double), cudaMemcpyDeviceToHost));
    return RT_OK;
}

extern "C" rt_status rt_set_gain_width_line(rt_model* m, int up, int low, double thermal_offset)
{
    if (!m) return RT_ERR_INVALID;
    if (up < 0) { m->gw_up = m->gw_low = -1; m->gw_offset = 0.; return RT_OK; }
    if (up >= m->d.N || low < 0 || low >= up) return RT_ERR_INVALID;
    m->gw_up = up; m->gw_low = low; m->gw_offset = thermal_offset;
    return RT_OK;
}

extern "C" rt_status rt_maser_postpass(rt_model* m, const double* pop, int nb_trans, const int* up, const int* low,
                                       int nb_freq, int nb_aspect, double delta_aspect, double velocity_shift,
                                       double* inv_arr, double* gain_arr, double* scalars, double* tau_vs_aspect,
                                       double* tau_vs_freq)
{
    NEED_CLOUD();
    const DevModel& d = m->d;
    const int L = d.L, N = d.N;
    if (!pop || nb_trans < 0 || nb_freq < 2 || nb_freq > 320 || nb_aspect < 1 || nb_aspect > 64) return RT_ERR_INVALID;
    if (nb_trans == 0) return RT_OK;
    rt_status r = rt_set_populations(m, pop);
    if (r) return r;
    std::vector<double> sig((size_t)nb_trans * d.ncomp), vgrid(nb_freq);
    for (int t = 0; t < nb_trans; t++) {
        if (up[t] <= low[t] || up[t] >= N || low[t] < 0) return RT_ERR_INVALID;
        const std::vector<double> s = line_sigma(m, m->h_energy[up[t]] - m->h_energy[low[t]]);
        std::copy(s.begin(), s.end(), sig.begin() + (size_t)t * d.ncomp);
    }
    // velocity grid, transition_data.cpp:312-315,343: vel += dv accumulated exactly like the reference's loop
    const double vmax = m->h_vel_n[0] + velocity_shift, vmin = m->h_vel_n[L - 1] - velocity_shift;
    const double dv = (vmax - vmin) / (nb_freq - 1.);
    double vel = vmin;
    for (int n = 0; n < nb_freq; n++, vel += dv) vgrid[n] = vel;
    DevScratch scratch(m->stream);
    bool ok = true;
    auto up_d = [&](const void* h, size_t bytes) { void* p = scratch.get(bytes, h); ok = ok && p; return p; };
    auto out_d = [&](size_t bytes) { void* p = scratch.get(bytes); ok = ok && p; return p; };
    PostArgs a{};
    a.pop = d.pop;
    a.up = (const int*)up_d(up, sizeof(int) * nb_trans); a.low = (const int*)up_d(low, sizeof(int) * nb_trans);
    a.sigma = (const double*)up_d(sig.data(), sizeof(double) * sig.size());
    a.dz = (const double*)up_d(m->h_dz.data(), sizeof(double) * L);
    a.vel_n = (const double*)up_d(m->h_vel_n.data(), sizeof(double) * L);
    a.vgrid = (const double*)up_d(vgrid.data(), sizeof(double) * nb_freq);
    a.vel_turb = (const double*)up_d(m->h_vel_turb.data(), sizeof(double) * L);
    a.gw_up = m->gw_up; a.gw_low = m->gw_low; a.gw_offset = m->gw_offset;
    a.height = m->h_height; a.delta_aspect = delta_aspect; a.nb_freq = nb_freq; a.nb_aspect = nb_aspect;
    a.inv_arr = (double*)out_d(sizeof(double) * (size_t)nb_trans * L);
    a.gain_arr = (double*)out_d(sizeof(double) * (size_t)nb_trans * L);
    a.scalars = (double*)out_d(sizeof(double) * 5 * (size_t)nb_trans);
    a.tau_aspect = (double*)out_d(sizeof(double) * (size_t)nb_trans * nb_aspect);
    a.tau_freq = (double*)out_d(sizeof(double) * (size_t)nb_trans * nb_freq);
    NEED_BUF(ok);
    const int c0 = rt_prof_class(m, "maser_postpass");
    rt_prof_begin(m, c0, m->stream);
    launch_postpass_impl(d, m->stream, a, nb_trans);
    rt_prof_end(m, c0, m->stream);
    CK(cudaStreamSynchronize(m->stream));
    CK(cudaGetLastError());
    if (inv_arr) CK(cudaMemcpy(inv_arr, a.inv_arr, sizeof(double) * (size_t)nb_trans * L, cudaMemcpyDeviceToHost));
    if (gain_arr) CK(cudaMemcpy(gain_arr, a.gain_arr, sizeof(double) * (size_t)nb_trans * L, cudaMemcpyDeviceToHost));
    if (scalars) CK(cudaMemcpy(scalars, a.scalars, sizeof(double) * 5 * (size_t)nb_trans, cudaMemcpyDeviceToHost));
    if (tau_vs_aspect) CK(cudaMemcpy(tau_vs_aspect, a.tau_aspect, sizeof(double) * (size_t)nb_trans * nb_aspect, cudaMemcpyDeviceToHost));
    if (tau_vs_freq) CK(cudaMemcpy(tau_vs_freq, a.tau_freq, sizeof(double) * (size_t)nb_trans * nb_freq, cudaMemcpyDeviceToHost));
    return RT_OK;
}

// ---- rt_lim_luminosity -------------------------------------------------------------------
void launch_loss_rates(const DevModel& d, cudaStream_t s, const double* pop_dev, const double* pop_intensity, size_t intensity_stride,
                       int nb_trans, const int* up_dev, const int* low_dev, double* out_dev);

extern "C" rt_status rt_lim_luminosity(rt_model* m, const double* pop, int nb_trans, const int* up, const int* low,
                                       int layer0_intensity, double* lum_arr, double* pump_eff_arr, double* loss_rate_arr,
                                       double* pump_rate_arr, double* emiss_coeff_arr, double* exc_temp_arr, double* lum,
                                       const int* lay_nb_hg, double beaming_factor, double* tau_sat, double* temp_sat)
{
    NEED_CLOUD();
    const DevModel& d = m->d;
    const int L = d.L, N = d.N;
    if (!pop || nb_trans < 0 || (nb_trans > 0 && (!up || !low))) return RT_ERR_INVALID;
    if (nb_trans == 0) return RT_OK;
    for (int t = 0; t < nb_trans; t++)
        if (low[t] < 0 || up[t] >= N || up[t] <= low[t]) { m->last_error = "transition must have up > low"; return RT_ERR_INVALID; }
    rt_status r = rt_set_populations(m, pop);
    if (r != RT_OK) return r;
    const size_t nout = (size_t)nb_trans * L * 2;
    DevScratch scratch(m->stream);
    int* up_dev = (int*)scratch.get(sizeof(int) * nb_trans, up);
    int* low_dev = (int*)scratch.get(sizeof(int) * nb_trans, low);
    double* out_dev = (double*)scratch.get(sizeof(double) * nout);
    NEED_BUF(up_dev && low_dev && out_dev);
    const int c_l = rt_prof_class(m, "loss_rates");
    m->prof_layers = L;
    rt_prof_begin(m, c_l, m->stream, 0., 0.);
    launch_loss_rates(d, m->stream, d.pop, d.pop, layer0_intensity ? 0 : (size_t)N, nb_trans, up_dev, low_dev, out_dev);
    rt_prof_end(m, c_l, m->stream);
    CK(cudaStreamSynchronize(m->stream));
    CK(cudaGetLastError());
    std::vector<double> rates(nout);
    CK(cudaMemcpy(rates.data(), out_dev, sizeof(double) * nout, cudaMemcpyDeviceToHost));

    // the scalar part in the reference's own expression order (maser_luminosity.cpp:78-101)
    const rt_params& pr = d.prm;
    for (int t = 0; t < nb_trans; t++) {
        const int u = up[t], lo = low[t];
        const double gu = m->h_g[u], gl = m->h_g[lo];
        const double energy = m->h_energy[u] - m->h_energy[lo];
        double lum_t = 0.;
        for (int l = 0; l < L; l++) {
            const size_t o = (size_t)t * L + l;
            const double low_loss = rates[2 * o], up_loss = rates[2 * o + 1];
            const double pu = pop[(size_t)l * N + u], pl = pop[(size_t)l * N + lo];
            const double inversion = pu / gu - pl / gl;
            double lum_l, eff_l;
            if (inversion > 0.) {
                lum_l = inversion / (1. / (up_loss * gu) + 1. / (low_loss * gl)) * m->h_mol_conc[l];
                eff_l = inversion / (pu / gu + pl / gl);
            }
            else lum_l = eff_l = 1.e-99;
            lum_t += lum_l * m->h_dz[l];
            if (lum_arr) lum_arr[o] = lum_l;
            if (pump_eff_arr) pump_eff_arr[o] = eff_l;
            if (emiss_coeff_arr) emiss_coeff_arr[o] = m->h_h2_conc[l] * m->h_mol_conc[l] / m->h_velg[l];
            if (loss_rate_arr) loss_rate_arr[o] = (up_loss * gu + low_loss * gl) / (gu + gl);
            if (pump_rate_arr) pump_rate_arr[o] = 0.5 * (pu * up_loss + pl * low_loss) / m->h_h2_conc[l];
            if (exc_temp_arr) exc_temp_arr[o] = pr.cm_inverse_to_kelvins * energy / log((pl * gu) / (pu * gl)); // transition_data.cpp:234
        }
        if (lum) lum[t] = lum_t / m->h_height;
        if (lay_nb_hg && (tau_sat || temp_sat)) { // transition_data.cpp:456-483
            const int l = lay_nb_hg[t];
            double ts = 0., tmp = 0.;
            if (l >= 0 && l < L) {
                const size_t o = (size_t)t * L + l;
                const double pu = pop[(size_t)l * N + u], pl = pop[(size_t)l * N + lo];
                const double exc_temp = pr.cm_inverse_to_kelvins * energy / log((pl * gu) / (pu * gl));
                if (exc_temp < 0.) {
                    const double source_function = 1. / (1. - exp(energy * pr.cm_inverse_to_kelvins / exc_temp));
                    const double cmb_intensity = 1. / (exp(energy * pr.cm_inverse_to_kelvins / 2.73) - 1.);
                    const double intensity_0 = source_function + cmb_intensity;
                    const double loss = (rates[2 * o + 1] * gu + rates[2 * o] * gl) / (gu + gl);
                    ts = log(loss / (m->h_einst[(size_t)u * N + lo] * beaming_factor * intensity_0));
                    tmp = intensity_0 * exp(ts) * pr.planck_constant * (pr.speed_of_light * energy) / pr.boltzmann_constant;
                }
            }
            if (tau_sat) tau_sat[t] = ts;
            if (temp_sat) temp_sat[t] = tmp;
        }
    }
    return RT_OK;
}
