/*
 * oracle/rt_oracle.c -- TEST INFRASTRUCTURE ONLY.  Plain-C CPU restatement of the level-population hot path of
 * nesterenok/radiative_transfer, taking the flat problem of include/rt_b200.h.  It is the checker for the CUDA
 * path (tests/, __graft_entry__.smoke(), bench.py's cpu_baseline "port" leg) and is never linked into, imported
 * by or called from the product (radiative_transfer_b200/).
 *
 * Parity status: the reference ships no tests or golden vectors.  This restatement is
 * pinned against the reference ITSELF: oracle/_ref/libref_oracle.so (the unmodified reference sources compiled
 * against oracle/shim/) on identical inputs -- tests/test_oracle_vs_reference.py -- and against the fixtures
 * under tests/golden/ generated from it (tests/golden/make_golden.py).  Two boundaries stay "parity unpinned"
 * because the reference omits them: lu_matrix_solve and the physical constants (the headers in oracle/shim/).
 *
 * Every function cites the reference file:line it follows (paths relative to radiative_transfer/).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#include <omp.h>
#include "../include/rt_b200.h"

/* ------------------------------------------------------------------------------------------------------- */
/* small helpers                                                                                           */
/* ------------------------------------------------------------------------------------------------------- */

/* oracle/shim/utils.h locate_index: arr[k] <= x < arr[k+1]; -1 below, n-1 at/above the last node */
static int locate_index(const double* arr, int n, double x)
{
    if (!(x >= arr[0])) return -1;
    if (x >= arr[n - 1]) return n - 1;
    int lo = 0, hi = n - 1;
    while (hi - lo > 1) {
        int mid = lo + ((hi - lo) >> 1);
        if (arr[mid] <= x) lo = mid; else hi = mid;
    }
    return lo;
}

/* collision_data::locate, coll_rates.cpp:71-82 (strict '<', never exceeds jmax-2) */
static int coll_locate(const double* tgrid, int jmax, double temp)
{
    int l = 0, r = jmax - 1;
    while (r - l > 1) {
        int j = l + ((r - l) >> 1);
        if (tgrid[j] < temp) l = j; else r = j;
    }
    return l;
}

/* Dense solve, same contract and pivot rule as oracle/shim/linear_algebra.h lu_matrix_solve (row partial
 * pivoting, first maximum wins).  a is [n][n] row-major, destroyed; b -> solution. */
static void lu_solve(double* a, double* b, int n)
{
    for (int k = 0; k < n; k++) {
        int p = k;
        double big = fabs(a[(size_t)k * n + k]);
        for (int i = k + 1; i < n; i++) {
            double v = fabs(a[(size_t)i * n + k]);
            if (v > big) { big = v; p = i; }
        }
        if (p != k) {
            double* rp = a + (size_t)p * n; double* rk = a + (size_t)k * n;
            for (int j = 0; j < n; j++) { double t = rp[j]; rp[j] = rk[j]; rk[j] = t; }
            double t = b[p]; b[p] = b[k]; b[k] = t;
        }
        const double piv = a[(size_t)k * n + k];
        if (piv == 0.) continue;
        const double rpiv = 1. / piv;
        const double* rk = a + (size_t)k * n;
        for (int i = k + 1; i < n; i++) {
            double* ri = a + (size_t)i * n;
            const double l = ri[k] * rpiv;
            if (l != 0.) {
                ri[k] = l;
                for (int j = k + 1; j < n; j++) ri[j] -= l * rk[j];
                b[i] -= l * b[k];
            }
        }
    }
    for (int i = n - 1; i >= 0; i--) {
        double s = b[i];
        const double* ri = a + (size_t)i * n;
        for (int j = i + 1; j < n; j++) s -= ri[j] * b[j];
        b[i] = s / ri[i];
    }
}

/* ------------------------------------------------------------------------------------------------------- */
/* model context                                                                                           */
/* ------------------------------------------------------------------------------------------------------- */
typedef struct ora_table {
    int nb_lev, jmax, imax, interp;
    const double* tgrid;
    const double* coeff;
    double* deriv; /* [imax][jmax], collision_data::calc_coeff_deriv, coll_rates.cpp:61-69; cubic spline (:89-119): second derivatives */
    double max_temp;
} ora_table;

typedef struct ora_dust {
    int n;
    const double* ph_en;
    const double* abs_coeff;
    double* deriv; /* dust_component::calc_derivatives, dust_model.cpp:516-524 */
    double wvl_exp;
} ora_dust;

typedef struct ora_overlap {
    int nb_d, nb_dx, nb_gr, nb_g;
    double* log_delta; /* lvg_method_functions.cpp:293 */
    const double *dx, *gratio, *gamma, *p;
} ora_overlap;

typedef struct ora_model {
    const rt_model_desc* d;
    int N, K, nb1, nb2;
    ora_table* tab;
    ora_dust* dust;
    int ndust;
    ora_overlap ov[2];
    int ngroups;
    int* groups; /* [ngroups][5] */
} ora_model;

static void overlap_init(ora_overlap* o, const rt_overlap_table* t)
{
    o->nb_d = t->nb_d; o->nb_dx = t->nb_dx; o->nb_gr = t->nb_gr; o->nb_g = t->nb_g;
    o->dx = t->dx; o->gratio = t->gratio; o->gamma = t->gamma; o->p = t->p;
    o->log_delta = (double*)malloc(sizeof(double) * t->nb_d);
    for (int i = 0; i < t->nb_d; i++) o->log_delta[i] = log10(t->delta[i]);
}

static int build_line_groups(const rt_model_desc* d, int** out);

static ora_model* model_create(const rt_model_desc* d)
{
    ora_model* m = (ora_model*)calloc(1, sizeof(ora_model));
    m->d = d;
    m->N = d->nb_lev;
    m->nb1 = d->nb_neutral_tables;
    m->nb2 = m->nb1 + d->nb_electron_tables;
    m->K = m->nb2;
    m->tab = (ora_table*)calloc(m->K > 0 ? m->K : 1, sizeof(ora_table));
    for (int k = 0; k < m->K; k++) {
        ora_table* t = &m->tab[k];
        const rt_coll_table* s = &d->tables[k];
        t->nb_lev = s->nb_lev; t->jmax = s->jmax; t->imax = s->nb_lev * (s->nb_lev - 1) / 2;
        t->tgrid = s->tgrid; t->coeff = s->coeff;
        t->max_temp = s->tgrid[s->jmax - 1]; /* get_max_temp, coll_rates.h:23 */
        t->deriv = (double*)calloc((size_t)t->imax * t->jmax + 1, sizeof(double));
        t->interp = s->interp;
        if (s->interp == RT_INTERP_CUBIC_SPLINE) {
            /* collision_data_cub_spline::calc_coeff_deriv, coll_rates.cpp:89-119 */
            double* u = (double*)calloc(t->jmax, sizeof(double));
            const double* tg = s->tgrid;
            for (int i = 0; i < t->imax; i++) {
                const double* c = s->coeff + (size_t)i * t->jmax;
                double* d2 = t->deriv + (size_t)i * t->jmax;
                d2[0] = u[0] = 0.;
                for (int j = 1; j < t->jmax - 1; j++) {
                    const double sig = (tg[j] - tg[j - 1]) / (tg[j + 1] - tg[j - 1]);
                    const double p = sig * d2[j - 1] + 2.;
                    d2[j] = (sig - 1.) / p;
                    u[j] = (c[j + 1] - c[j]) / (tg[j + 1] - tg[j]) - (c[j] - c[j - 1]) / (tg[j] - tg[j - 1]);
                    u[j] = (6. * u[j] / (tg[j + 1] - tg[j - 1]) - sig * u[j - 1]) / p;
                }
                d2[t->jmax - 1] = 0.;
                for (int j = t->jmax - 2; j >= 0; j--) d2[j] = d2[j] * d2[j + 1] + u[j];
            }
            free(u);
        }
        else
            for (int i = 0; i < t->imax; i++)
                for (int j = 0; j < t->jmax - 1; j++)
                    t->deriv[(size_t)i * t->jmax + j] = (s->coeff[(size_t)i * t->jmax + j + 1] - s->coeff[(size_t)i * t->jmax + j])
                                                        / (s->tgrid[j + 1] - s->tgrid[j]);
    }
    m->ndust = d->dust_ncomp;
    m->dust = (ora_dust*)calloc(m->ndust > 0 ? m->ndust : 1, sizeof(ora_dust));
    for (int c = 0; c < m->ndust; c++) {
        ora_dust* u = &m->dust[c];
        u->n = d->dust[c].nb_ph_en; u->ph_en = d->dust[c].ph_en; u->abs_coeff = d->dust[c].abs_coeff;
        u->wvl_exp = d->dust[c].wvl_exp;
        u->deriv = (double*)calloc(u->n, sizeof(double));
        for (int i = 0; i < u->n - 1; i++)
            u->deriv[i] = (u->abs_coeff[i + 1] - u->abs_coeff[i]) / (u->ph_en[i + 1] - u->ph_en[i]);
    }
    if (d->scheme == RT_SCHEME_LINE_OVERLAP) {
        overlap_init(&m->ov[0], d->overlap1);
        overlap_init(&m->ov[1], d->overlap2);
        if (d->nb_line_groups > 0) {
            m->ngroups = d->nb_line_groups;
            m->groups = (int*)malloc(sizeof(int) * 5 * m->ngroups);
            memcpy(m->groups, d->line_groups, sizeof(int) * 5 * m->ngroups);
        }
        else m->ngroups = build_line_groups(d, &m->groups);
    }
    return m;
}

static void model_destroy(ora_model* m)
{
    for (int k = 0; k < m->K; k++) free(m->tab[k].deriv);
    free(m->tab);
    for (int c = 0; c < m->ndust; c++) free(m->dust[c].deriv);
    free(m->dust);
    free(m->ov[0].log_delta); free(m->ov[1].log_delta);
    free(m->groups);
    free(m);
}

/* ------------------------------------------------------------------------------------------------------- */
/* OH hyperfine line groups: iteration_scheme_line_overlap::init_molecule_data, iteration_lvg.cpp:317-346,  */
/* with hfs_lines::sort (:259-290) and split (:292-302), quirks kept (the "closest pair" search at          */
/* :274-280 compares against a never-updated `e`).                                                          */
/* ------------------------------------------------------------------------------------------------------- */
typedef struct hfs { int nb; int up[4], low[4]; double en[4]; } hfs;

static void hfs_swap(hfs* h, int i, int j)
{
    int t = h->up[i]; h->up[i] = h->up[j]; h->up[j] = t;
    t = h->low[i]; h->low[i] = h->low[j]; h->low[j] = t;
    double e = h->en[i]; h->en[i] = h->en[j]; h->en[j] = e;
}

static void hfs_sort(hfs* h)
{
    if (h->nb <= 2) return;
    int i, j;
    for (i = 0; i < h->nb; i++)
        for (j = i + 1; j < h->nb; j++)
            if (h->en[j] < h->en[i]) hfs_swap(h, i, j);
    j = 0;
    double e = h->en[1] - h->en[0];
    for (i = 1; i < h->nb - 1; i++)
        if (h->en[i + 1] - h->en[i] < e) j = i; /* e is not updated: reference behaviour */
    e = h->en[j];
    for (i = 0; i < h->nb; i++)
        for (j = i + 1; j < h->nb; j++)
            if (fabs(h->en[j] - e) < fabs(h->en[i] - e)) hfs_swap(h, i, j);
}

static int build_line_groups(const rt_model_desc* d, int** out)
{
    const int N = d->nb_lev;
    int cap = 16, n = 0;
    int* g = (int*)malloc(sizeof(int) * 5 * cap);
    for (int i = 2; i < N; i += 2)
        for (int j = 0; j < i; j += 2) {
            hfs h; h.nb = 0;
            for (int m = 0; m < 2; m++)
                for (int l = 0; l < 2; l++) {
                    /* the reference indexes arr[i+m][j+l] without a bound check; for odd N the last block would
                     * read out of range there -- guarded here (i + m < N) */
                    if (i + m < N && d->einst[(size_t)(i + m) * N + j + l] > 1.e-99) {
                        h.up[h.nb] = i + m; h.low[h.nb] = j + l;
                        h.en[h.nb] = d->energy[i + m] - d->energy[j + l];
                        h.nb++;
                    }
                }
            hfs_sort(&h);
            if (n + 2 > cap) { cap *= 2; g = (int*)realloc(g, sizeof(int) * 5 * cap); }
            int first = 0;
            if (h.nb >= 3) { /* split(): the first two lines become their own group */
                int* q = g + 5 * n++;
                q[0] = 2; q[1] = h.up[0]; q[2] = h.low[0]; q[3] = h.up[1]; q[4] = h.low[1];
                first = 2;
            }
            int rest = h.nb - first;
            if (rest > 0) {
                int* q = g + 5 * n++;
                q[0] = rest; q[1] = h.up[first]; q[2] = h.low[first];
                q[3] = rest > 1 ? h.up[first + 1] : -1; q[4] = rest > 1 ? h.low[first + 1] : -1;
            }
        }
    *out = g;
    return n;
}

/* ------------------------------------------------------------------------------------------------------- */
/* per-layer state: the three setters of radiative_transfer.cpp:238-242                                     */
/* ------------------------------------------------------------------------------------------------------- */
#define ORA_MAXK 16
typedef struct ora_layer {
    double temp_n, temp_el, mol_conc, vel_width, vel_grad;
    double conc[ORA_MAXK];
    int idx[ORA_MAXK];
    const double* dust_conc;
} ora_layer;

/* collisional_transitions::set_gas_param coll_rates.cpp:152-174 + overrides coll_rates_ch3oh.cpp:472-481,
 * coll_rates_oh.cpp:380-389, coll_rates_h2o.cpp:515-527 */
static void set_gas_param(const ora_model* m, ora_layer* s, double tn, double te, double he, double ph2, double oh2,
                          double h, double el)
{
    for (int k = 0; k < m->K; k++) s->conc[k] = 0.;
    for (int k = 0; k < m->nb1; k++) s->idx[k] = coll_locate(m->tab[k].tgrid, m->tab[k].jmax, tn);
    for (int k = m->nb1; k < m->nb2; k++) {
        s->conc[k] = el;
        s->idx[k] = coll_locate(m->tab[k].tgrid, m->tab[k].jmax, te);
    }
    switch (m->d->species_rule) {
    case RT_RULE_CH3OH:
    case RT_RULE_OH_HF:
    case RT_RULE_H2CO: /* coll_rates_h2co.cpp:160-168 */
        s->conc[0] = he; s->conc[1] = ph2; s->conc[2] = oh2;
        break;
    case RT_RULE_NH3: /* coll_rates_nh3.cpp:261-271 */
        s->conc[0] = he; s->conc[1] = ph2; s->conc[2] = oh2; s->conc[3] = h;
        break;
    case RT_RULE_H2O:
        s->conc[0] = he; s->conc[1] = he + 0.2 * h; s->conc[2] = ph2; s->conc[3] = oh2; s->conc[4] = ph2 + oh2; s->conc[5] = h;
        break;
    default: { /* generic convention of this build: partners in the order He, p-H2, o-H2, H */
        const double c[4] = {he, ph2, oh2, h};
        for (int k = 0; k < m->nb1; k++) s->conc[k] = c[k < 4 ? k : 3];
    }
    }
}

static void layer_set(const ora_model* m, const rt_layer_soa* L, int l, ora_layer* s)
{
    const rt_params* P = &m->d->params;
    s->temp_n = L->temp_n[l];
    s->temp_el = L->temp_el[l];
    s->mol_conc = L->mol_conc[l];
    s->vel_grad = L->velg_n[l];
    /* iteration_lvg.cpp:65 */
    s->vel_width = pow(2. * P->boltzmann_constant * s->temp_n / m->d->mol_mass + L->vel_turb[l] * L->vel_turb[l], 0.5);
    s->dust_conc = L->dust_conc + (size_t)l * L->dust_ncomp;
    set_gas_param(m, s, s->temp_n, s->temp_el, L->he_conc[l], L->ph2_conc[l], L->oh2_conc[l], L->h_conc[l], L->el_conc[l]);
}

/* collision_data::get_rate(first, sec, lo, temp) coll_rates.cpp:54-59 with the caller-side clamp of temp to the
 * table maximum (coll_rates.cpp:188) */
static double tab_rate(const ora_table* t, int up, int low, int lo, double temp)
{
    const size_t i = (size_t)up * (up - 1) / 2 + low;
    const double tt = temp < t->max_temp ? temp : t->max_temp;
    if (t->interp == RT_INTERP_CUBIC_SPLINE) { /* collision_data_cub_spline::get_rate, coll_rates.cpp:121-132 */
        const double h = t->tgrid[lo + 1] - t->tgrid[lo];
        const double a = (t->tgrid[lo + 1] - tt) / h, b = 1. - a;
        const double* c = t->coeff + i * t->jmax;
        const double* s2 = t->deriv + i * t->jmax;
        return fabs(a * c[lo] + b * c[lo + 1] + (1. / 6.) * ((a * a * a - a) * s2[lo] + (b * b * b - b) * s2[lo + 1]) * (h * h));
    }
    return t->coeff[i * t->jmax + lo] + t->deriv[i * t->jmax + lo] * (tt - t->tgrid[lo]);
}

/* get_rate_neutrals: base coll_rates.cpp:181-197; ch3oh coll_rates_ch3oh.cpp:484-533 (USE_TEMPER_EXTRAP_CH3OH 0);
 * oh_hf coll_rates_oh.cpp:392-407; h2o coll_rates_h2o.cpp:530-548 */
static void rate_neutrals(const ora_model* m, const ora_layer* s, int up, int low, double* down, double* upr)
{
    const rt_model_desc* d = m->d;
    const ora_table* T = m->tab;
    const double tn = s->temp_n;
    double dn = 0.;
    switch (d->species_rule) {
    case RT_RULE_CH3OH:
        if (d->v[up] == d->v[low]) {
            if (d->v[up] == 0 && d->j[up] <= 9 && d->j[low] <= 9)
                dn = tab_rate(&T[1], up, low, s->idx[1], tn) * s->conc[1] + tab_rate(&T[2], up, low, s->idx[2], tn) * s->conc[2];
            else
                dn = tab_rate(&T[1], up, low, s->idx[1], tn) * (s->conc[1] + s->conc[2]);
            dn += tab_rate(&T[0], up, low, s->idx[0], tn) * s->conc[0];
        }
        else dn = tab_rate(&T[0], up, low, s->idx[0], tn) * (s->conc[0] + s->conc[1] + 3. * s->conc[2]);
        break;
    case RT_RULE_OH_HF:
        if (up < T[0].nb_lev) dn = tab_rate(&T[0], up, low, s->idx[0], tn) * s->conc[0];
        if (up < T[1].nb_lev)
            dn += tab_rate(&T[1], up, low, s->idx[1], tn) * s->conc[1] + tab_rate(&T[2], up, low, s->idx[2], tn) * s->conc[2];
        break;
    case RT_RULE_H2O:
        if (up < 45)
            dn = tab_rate(&T[0], up, low, s->idx[0], tn) * s->conc[0] + tab_rate(&T[2], up, low, s->idx[2], tn) * s->conc[2]
                 + tab_rate(&T[3], up, low, s->idx[3], tn) * s->conc[3] + tab_rate(&T[5], up, low, s->idx[5], tn) * s->conc[5];
        else
            dn = tab_rate(&T[1], up, low, s->idx[1], tn) * s->conc[1] + tab_rate(&T[4], up, low, s->idx[4], tn) * s->conc[4];
        break;
    case RT_RULE_NH3: /* coll_rates_nh3.cpp:273-290 */
        if (up < T[0].nb_lev) dn = tab_rate(&T[0], up, low, s->idx[0], tn) * s->conc[0];
        if (up < T[1].nb_lev)
            dn += tab_rate(&T[1], up, low, s->idx[1], tn) * s->conc[1] + tab_rate(&T[2], up, low, s->idx[2], tn) * s->conc[2]
                  + tab_rate(&T[3], up, low, s->idx[3], tn) * s->conc[3];
        break;
    case RT_RULE_H2CO: /* coll_rates_h2co.cpp:170-185 */
        if (up < T[0].nb_lev) dn = tab_rate(&T[0], up, low, s->idx[0], tn) * s->conc[0];
        if (up < T[1].nb_lev)
            dn += tab_rate(&T[1], up, low, s->idx[1], tn) * s->conc[1] + tab_rate(&T[2], up, low, s->idx[2], tn) * s->conc[2];
        break;
    default:
        for (int k = 0; k < m->nb1; k++)
            if (up < T[k].nb_lev) dn += tab_rate(&T[k], up, low, s->idx[k], tn) * s->conc[k];
    }
    if (dn > d->params.min_collision_rate) {
        *down = dn;
        *upr = dn * exp((d->energy[low] - d->energy[up]) * d->params.cm_inverse_to_kelvins / tn) * d->g[up] / ((double)d->g[low]);
    }
    else *down = *upr = 0.;
}

/* get_rate_electrons coll_rates.cpp:199-217: only the FIRST applicable table */
static void rate_electrons(const ora_model* m, const ora_layer* s, int up, int low, double* down, double* upr)
{
    const rt_model_desc* d = m->d;
    double dn = 0.;
    for (int k = m->nb1; k < m->nb2; k++)
        if (up < m->tab[k].nb_lev) {
            dn = tab_rate(&m->tab[k], up, low, s->idx[k], s->temp_el) * s->conc[k];
            break;
        }
    if (dn > d->params.min_collision_rate) {
        *down = dn;
        *upr = dn * exp((d->energy[low] - d->energy[up]) * d->params.cm_inverse_to_kelvins / s->temp_el) * d->g[up] / ((double)d->g[low]);
    }
    else *down = *upr = 0.;
}

/* ------------------------------------------------------------------------------------------------------- */
/* table lookups                                                                                           */
/* ------------------------------------------------------------------------------------------------------- */

/* lvg_method_data::get_esc_func, lvg_method_functions.cpp:74-110 */
static double esc_func(const rt_lvg_table* t, double gamma, double delta)
{
    int k = locate_index(t->delta, t->nb_d, delta);
    int l = locate_index(t->gamma, t->nb_g, gamma);
    double tt, u;
    if (k < 0) { tt = 0.; k = 0; }
    else if (k > t->nb_d - 2) { tt = 1.; k = t->nb_d - 2; }
    else tt = (delta - t->delta[k]) / (t->delta[k + 1] - t->delta[k]);
    if (l < 0) { l = 0; u = 0.; }
    else if (l > t->nb_g - 2) { l = t->nb_g - 2; u = 1.; }
    else u = (gamma - t->gamma[l]) / (t->gamma[l + 1] - t->gamma[l]);
    const double* p = t->p;
    const int ng = t->nb_g;
    double e = p[k * ng + l] * (1. - tt) * (1. - u) + p[(k + 1) * ng + l] * tt * (1. - u) + p[k * ng + l + 1] * (1. - tt) * u
               + p[(k + 1) * ng + l + 1] * u * tt;
    if (e > 1.) e = 1.;
    else if (e < 0.) e = 0.;
    return e;
}

/* lvg_line_overlap_data::get_esc_func, lvg_method_functions.cpp:324-392 (term order kept) */
static double overlap_esc_func(const ora_overlap* o, double gamma, double delta, double gamma_ratio, double delta_x)
{
    delta = log10(delta);
    int m = locate_index(o->log_delta, o->nb_d, delta);
    int l = locate_index(o->gamma, o->nb_g, gamma);
    int k = locate_index(o->gratio, o->nb_gr, gamma_ratio);
    int n = locate_index(o->dx, o->nb_dx, delta_x);
    double y = 0., u = 0., t = 0., p = 0.;
    if (m < 0) m = 0;
    else if (m > o->nb_d - 2) { m = o->nb_d - 2; y = 1.; }
    else y = (delta - o->log_delta[m]) / (o->log_delta[m + 1] - o->log_delta[m]);
    if (n < 0) n = 0;
    else if (n > o->nb_dx - 2) { p = 1.; n = o->nb_dx - 2; }
    else p = (delta_x - o->dx[n]) / (o->dx[n + 1] - o->dx[n]);
    if (l < 0) l = 0;
    else if (l > o->nb_g - 2) { l = o->nb_g - 2; u = 1.; }
    else u = (gamma - o->gamma[l]) / (o->gamma[l + 1] - o->gamma[l]);
    if (k < 0) k = 0;
    else if (k > o->nb_gr - 2) { t = 1.; k = o->nb_gr - 2; }
    else t = (gamma_ratio - o->gratio[k]) / (o->gratio[k + 1] - o->gratio[k]);
    const size_t W = (size_t)o->nb_g * o->nb_gr;
    const int ndx = o->nb_dx, ng = o->nb_g;
#define PA(a, b, c, e) o->p[(size_t)((m + (a)) * ndx + n + (b)) * W + (size_t)(k + (c)) * ng + l + (e)]
    double escf = PA(0, 0, 0, 0) * (1. - u) * (1. - t) * (1. - p) * (1. - y)
                  + PA(0, 0, 0, 1) * u * (1. - t) * (1. - p) * (1. - y)
                  + PA(0, 0, 1, 0) * (1. - u) * t * (1. - p) * (1. - y)
                  + PA(0, 0, 1, 1) * u * t * (1. - p) * (1. - y)
                  + PA(0, 1, 0, 0) * (1. - u) * (1. - t) * p * (1. - y)
                  + PA(0, 1, 0, 1) * u * (1. - t) * p * (1. - y)
                  + PA(0, 1, 1, 0) * (1. - u) * t * p * (1. - y)
                  + PA(0, 1, 1, 1) * u * t * p * (1. - y)
                  + PA(1, 0, 0, 0) * (1. - u) * (1. - t) * (1. - p) * y
                  + PA(1, 0, 0, 1) * u * (1. - t) * (1. - p) * y
                  + PA(1, 0, 1, 0) * (1. - u) * t * (1. - p) * y
                  + PA(1, 0, 1, 1) * u * t * (1. - p) * y
                  + PA(1, 1, 0, 0) * (1. - u) * (1. - t) * p * y
                  + PA(1, 1, 0, 1) * u * (1. - t) * p * y
                  + PA(1, 1, 1, 0) * (1. - u) * t * p * y
                  + PA(1, 1, 1, 1) * u * t * p * y;
#undef PA
    if (escf > 1.) escf = 1.;
    else if (escf < 0.) escf = 0.;
    return escf;
}

/* dust_component::absorption dust_model.cpp:473-490; dust_model::absorption :834-841 */
static double dust_comp_absorption(const ora_dust* u, double energy)
{
    if (energy < u->ph_en[0]) return u->abs_coeff[0] * pow(u->ph_en[0] / energy, u->wvl_exp);
    if (energy > u->ph_en[u->n - 1]) return u->abs_coeff[u->n - 1];
    int l = 0, r = u->n - 1;
    while (r - l > 1) {
        int i = l + ((r - l) >> 1);
        if (u->ph_en[i] < energy) l = i; else r = i;
    }
    return u->abs_coeff[l] + u->deriv[l] * (energy - u->ph_en[l]);
}

static double dust_absorption(const ora_model* m, double energy, const double* conc)
{
    double a = 0.;
    for (int c = 0; c < m->ndust; c++) a += dust_comp_absorption(&m->dust[c], energy) * conc[c];
    return a;
}

/* ------------------------------------------------------------------------------------------------------- */
/* mean intensities                                                                                        */
/* ------------------------------------------------------------------------------------------------------- */

/* iteration_scheme_lvg::intensity_calc, iteration_lvg.cpp:163-185 */
static double intensity_single(const ora_model* m, const ora_layer* s, int up, int low, const double* pop)
{
    const rt_model_desc* d = m->d;
    const int N = m->N;
    const double energy = d->energy[up] - d->energy[low];
    const double c = s->mol_conc / (d->params.eight_pi * s->vel_width * energy * energy * energy);
    const double line_emiss = c * d->einst[(size_t)up * N + low] * pop[up];
    double line_opacity = c * d->einst[(size_t)low * N + up] * pop[low] - line_emiss + d->params.min_line_opacity;
    if (line_opacity < 0.) line_opacity *= d->params.inv_trans_factor;
    const double dust_opacity = dust_absorption(m, energy, s->dust_conc);
    const double gamma = fabs(s->vel_grad) / (s->vel_width * line_opacity);
    const double delta = fabs(s->vel_grad) / (s->vel_width * dust_opacity);
    const double ep1 = esc_func(&d->lvg, gamma, delta);
    return line_emiss / line_opacity * ep1;
}

/* iteration_scheme_line_overlap::intensity_calc(u1,l1,u2,l2,...), iteration_lvg.cpp:428-501 */
static void intensity_pair(const ora_model* m, const ora_layer* s, int u1, int l1, int u2, int l2, const double* pop,
                           double* intens1, double* intens2)
{
    const rt_model_desc* d = m->d;
    const int N = m->N;
    const double max_dx = d->params.max_dx;
    const double energy = d->energy[u1] - d->energy[l1];
    double c = s->mol_conc / (d->params.eight_pi * s->vel_width * energy * energy * energy);
    const double A1ul = d->einst[(size_t)u1 * N + l1], A1lu = d->einst[(size_t)l1 * N + u1];
    const double A2ul = d->einst[(size_t)u2 * N + l2], A2lu = d->einst[(size_t)l2 * N + u2];
    const double line_emiss1 = c * A1ul * pop[u1];
    double line_opacity1 = c * (A1lu * pop[l1] - A1ul * pop[u1]) + d->params.min_line_opacity;
    const double line_emiss2 = c * A2ul * pop[u2];
    double line_opacity2 = c * (A2lu * pop[l2] - A2ul * pop[u2]) + d->params.min_line_opacity;
    if (line_opacity1 < 0.) line_opacity1 *= d->params.inv_trans_factor;
    if (line_opacity2 < 0.) line_opacity2 *= d->params.inv_trans_factor;
    const double gamma1 = fabs(s->vel_grad) / (s->vel_width * line_opacity1);
    const double gamma2 = fabs(s->vel_grad) / (s->vel_width * line_opacity2);
    const double dust_opacity = dust_absorption(m, energy, s->dust_conc);
    const double delta = fabs(s->vel_grad) / (s->vel_width * dust_opacity);
    double dx = (d->energy[u1] - d->energy[l1] - d->energy[u2] + d->energy[l2]) * d->params.speed_of_light / (energy * s->vel_width);
    if (s->vel_grad < 0.) dx *= -1.;
    double ep1 = 0., ep2 = 0., ep01 = 0., ep02 = 0.;
    if (fabs(dx) < max_dx) {
        ep1 = overlap_esc_func(&m->ov[0], gamma1, delta, gamma2 / gamma1, dx);
        ep2 = overlap_esc_func(&m->ov[0], gamma2, delta, gamma1 / gamma2, -dx);
    }
    if (fabs(dx) > max_dx - 0.5) {
        ep01 = esc_func(&d->lvg, gamma1, delta);
        ep02 = esc_func(&d->lvg, gamma2, delta);
    }
    if (fabs(dx) > max_dx) { ep1 = ep01; ep2 = ep02; }
    else if (fabs(dx) > max_dx - 0.5) {
        c = 2. * (max_dx - fabs(dx));
        ep1 = ep01 * (1. - c) + ep1 * c;
        ep2 = ep02 * (1. - c) + ep2 * c;
    }
    *intens1 = line_emiss1 / line_opacity1 * ep1;
    *intens2 = line_emiss2 / line_opacity2 * ep2;
    if (fabs(dx) < max_dx) {
        ep1 = overlap_esc_func(&m->ov[1], gamma1, delta, gamma2 / gamma1, dx);
        ep2 = overlap_esc_func(&m->ov[1], gamma2, delta, gamma1 / gamma2, -dx);
        if (fabs(dx) > max_dx - 0.5) {
            c = 2. * (max_dx - fabs(dx));
            ep1 *= c;
            ep2 *= c;
        }
        *intens1 += line_emiss2 / line_opacity2 * ep1;
        *intens2 += line_emiss1 / line_opacity1 * ep2;
    }
}

/* virtual intensity_calc(up, low, ...) of the active scheme; overlap: iteration_lvg.cpp:503-526 (0 when the
 * line is not in line_list) */
static double intensity_any(const ora_model* m, const ora_layer* s, int up, int low, const double* pop)
{
    if (m->d->scheme != RT_SCHEME_LINE_OVERLAP) return intensity_single(m, s, up, low, pop);
    for (int i = 0; i < m->ngroups; i++) {
        const int* g = m->groups + 5 * i;
        for (int k = 0; k < g[0]; k++)
            if (g[2 + 2 * k] == low && g[1 + 2 * k] == up) {
                if (g[0] == 1) return intensity_single(m, s, up, low, pop);
                double a, b;
                intensity_pair(m, s, g[1], g[2], g[3], g[4], pop, &a, &b);
                return k == 0 ? a : b;
            }
    }
    return 0.;
}

/* ------------------------------------------------------------------------------------------------------- */
/* matrix assembly, step, start vectors                                                                    */
/* ------------------------------------------------------------------------------------------------------- */
static void add_line(double* M, int N, int u, int l, double Aul, double Alu, double J)
{
    /* iteration_lvg.cpp:137-144 */
    double y = Aul * (1. + J);
    M[(size_t)u * N + u] -= y;
    M[(size_t)l * N + u] += y;
    y = Alu * J;
    M[(size_t)u * N + l] += y;
    M[(size_t)l * N + l] -= y;
}

/* iteration_scheme_lvg::operator() iteration_lvg.cpp:112-161; line-overlap variant :348-426 */
static void assemble(const ora_model* m, const ora_layer* s, const double* pop, double* M, double* df)
{
    const rt_model_desc* d = m->d;
    const int N = m->N;
    const int overlap = d->scheme == RT_SCHEME_LINE_OVERLAP;
    memset(M, 0, sizeof(double) * (size_t)N * N);
    for (int i = 1; i < N; i++)
        for (int j = 0; j < i; j++) {
            double down, up, a, b;
            rate_neutrals(m, s, i, j, &down, &up);
            rate_electrons(m, s, i, j, &a, &b);
            down += a;
            up += b;
            M[(size_t)i * N + i] -= down;
            M[(size_t)j * N + i] += down;
            M[(size_t)i * N + j] += up;
            M[(size_t)j * N + j] -= up;
            if (!overlap && d->einst[(size_t)i * N + j] > 0.) {
                const double J = intensity_single(m, s, i, j, pop);
                add_line(M, N, i, j, d->einst[(size_t)i * N + j], d->einst[(size_t)j * N + i], J);
            }
        }
    if (overlap)
        for (int i = 0; i < m->ngroups; i++) {
            const int* g = m->groups + 5 * i;
            if (g[0] == 1) {
                const double J = intensity_single(m, s, g[1], g[2], pop);
                add_line(M, N, g[1], g[2], d->einst[(size_t)g[1] * N + g[2]], d->einst[(size_t)g[2] * N + g[1]], J);
            }
            else if (g[0] == 2) {
                double J1, J2;
                intensity_pair(m, s, g[1], g[2], g[3], g[4], pop, &J1, &J2);
                add_line(M, N, g[1], g[2], d->einst[(size_t)g[1] * N + g[2]], d->einst[(size_t)g[2] * N + g[1]], J1);
                add_line(M, N, g[3], g[4], d->einst[(size_t)g[3] * N + g[4]], d->einst[(size_t)g[4] * N + g[3]], J2);
            }
        }
    for (int j = 0; j < N; j++) M[j] = 1.;
    if (df) {
        memset(df, 0, sizeof(double) * N);
        df[0] = 1.;
        for (int i = 0; i < N; i++)
            for (int j = 0; j < N; j++) df[i] -= M[(size_t)i * N + j] * pop[j];
    }
}

/* iteration_scheme_lvg::calc_new_pop iteration_lvg.cpp:87-110; work = [N*N + N] scratch */
static void calc_new_pop(const ora_model* m, const ora_layer* s, const double* old_pop, double* new_pop, double* eq_error,
                         double* work)
{
    const int N = m->N;
    double* M = work;
    double* f = work + (size_t)N * N;
    assemble(m, s, old_pop, M, f);
    memset(new_pop, 0, sizeof(double) * N);
    new_pop[0] = 1.;
    lu_solve(M, new_pop, N);
    double e = 0.;
    for (int i = 0; i < N; i++)
        if (e < fabs(f[i])) e = fabs(f[i]);
    *eq_error = e;
}

/* boundary_layer_populations iteration_control.cpp:52-91 (neutrals only, 0.5*A, no J) */
static void boundary_populations(const ora_model* m, const ora_layer* s, double* arr, double* work)
{
    const rt_model_desc* d = m->d;
    const int N = m->N;
    double* M = work;
    memset(M, 0, sizeof(double) * (size_t)N * N);
    memset(arr, 0, sizeof(double) * N);
    for (int i = 1; i < N; i++)
        for (int j = 0; j < i; j++) {
            double down, up;
            rate_neutrals(m, s, i, j, &down, &up);
            const double a = 0.5 * d->einst[(size_t)i * N + j] + down;
            M[(size_t)j * N + i] = a;
            M[(size_t)i * N + i] -= a;
            M[(size_t)i * N + j] = up;
            M[(size_t)j * N + j] -= up;
        }
    for (int i = 0; i < N; i++) M[i] = 1.;
    arr[0] = 1.;
    lu_solve(M, arr, N);
}

/* boltzmann_populations spectroscopy.cpp:1245-1259 */
static void boltzmann(const ora_model* m, double temp, double* arr)
{
    const rt_model_desc* d = m->d;
    double sum = 0.;
    for (int i = 0; i < m->N; i++) {
        arr[i] = d->g[i] * exp(-d->energy[i] * d->params.cm_inverse_to_kelvins / temp);
        sum += arr[i];
    }
    for (int i = 0; i < m->N; i++) arr[i] /= sum;
}

/* ------------------------------------------------------------------------------------------------------- */
/* iteration control: iteration_control<T>, iteration_control.h:69-242                                      */
/* History = ring of nb_prev_steps+1 = 6 population and 6 residual vectors; slot 0 is "front".             */
/* ------------------------------------------------------------------------------------------------------- */
#define ORA_HIST 6
typedef struct ora_ctl {
    int N, n_pop, n_res;       /* list sizes */
    double* pop[ORA_HIST];     /* prev_level_pop, [0] = front */
    double* res[ORA_HIST];     /* residual_list,  [0] = front */
    double* opt;
    double* buf;
} ora_ctl;

static void ctl_init(ora_ctl* c, int N)
{
    c->N = N;
    c->buf = (double*)malloc(sizeof(double) * (size_t)N * (2 * ORA_HIST + 1));
    for (int i = 0; i < ORA_HIST; i++) { c->pop[i] = c->buf + (size_t)i * N; c->res[i] = c->buf + (size_t)(ORA_HIST + i) * N; }
    c->opt = c->buf + (size_t)2 * ORA_HIST * N;
}

static void push_front(double** ring, int* n, const double* v, int N)
{
    double* last = ring[ORA_HIST - 1];
    for (int i = ORA_HIST - 1; i > 0; i--) ring[i] = ring[i - 1];
    ring[0] = last;
    memcpy(last, v, sizeof(double) * N);
    if (*n < ORA_HIST) (*n)++; /* == pop_back when size > nb_prev_steps + 1, iteration_control.h:122-126 */
}

/* accel_step iteration_control.h:139-193, nb_prev_steps = 5 -> 4 unknowns */
static void accel_step(ora_ctl* c, double* accel_pop)
{
    const int N = c->N, np = 4;
    double A[16], b[4];
    memset(A, 0, sizeof(A));
    memset(b, 0, sizeof(b));
    const double* r0 = c->res[0];
    const double* n0 = c->pop[0];
    for (int i = 0; i < np; i++) {
        for (int j = 0; j < np; j++)
            for (int k = 0; k < N; k++) {
                const double w = n0[k] + 1.e-99;
                A[i * np + j] += (r0[k] - c->res[i + 1][k]) * (r0[k] - c->res[j + 1][k]) / (w * w);
            }
        for (int k = 0; k < N; k++) {
            const double w = n0[k] + 1.e-99;
            b[i] += (r0[k] - c->res[i + 1][k]) * r0[k] / (w * w);
        }
    }
    lu_solve(A, b, np);
    double sum = 0.;
    for (int i = 0; i < np; i++) sum += b[i];
    for (int k = 0; k < N; k++) {
        double v = (1. - sum) * c->pop[0][k];
        for (int i = 0; i < np; i++) v += b[i] * c->pop[i + 1][k];
        accel_pop[k] = v;
    }
}

/* calculate_populations iteration_control.h:196-242 with next_step :84-137 inlined */
static int calculate_populations(const ora_model* m, const ora_layer* s, double* pop, int max_iter, double min_error,
                                 int accel, rt_layer_status* st, double* work, ora_ctl* c)
{
    const rt_params* P = &m->d->params;
    const int N = m->N;
    int iter_nb = 0, nb_after_accel = 0, found = 0;
    double best_eq_error = 1., eq_error = 0., pop_error = 0., rel_error = 0.;
    double* pop_new = work + (size_t)N * N + N;
    double* lu_work = work;
    c->n_pop = c->n_res = 0;
    memset(c->opt, 0, sizeof(double) * N);
    do {
        push_front(c->pop, &c->n_pop, pop, N);
        if (accel && (iter_nb == P->accel_start || nb_after_accel == P->accel_period)) {
            accel_step(c, pop);
            nb_after_accel = 0;
        }
        calc_new_pop(m, s, pop, pop_new, &eq_error, lu_work);
        if (accel && iter_nb >= P->accel_start) nb_after_accel++;
        if (eq_error < best_eq_error) {
            best_eq_error = eq_error;
            memcpy(c->opt, pop, sizeof(double) * N);
        }
        pop_error = rel_error = 0.;
        double* r = c->res[ORA_HIST - 1]; /* slot that push_front recycles next */
        for (int i = 0; i < N; i++) {
            r[i] = pop_new[i] - pop[i];
            double a = fabs(r[i]);
            if (a > pop_error) pop_error = a;
            a = fabs(r[i] / (pop[i] + 1.e-99));
            if (a > rel_error) rel_error = a;
        }
        {   /* residual_list.push_front(residual): rotate the ring, data already in the recycled slot */
            double* last = c->res[ORA_HIST - 1];
            for (int i = ORA_HIST - 1; i > 0; i--) c->res[i] = c->res[i - 1];
            c->res[0] = last;
            if (c->n_res < ORA_HIST) c->n_res++;
        }
        if (iter_nb < max_iter - 1) memcpy(pop, pop_new, sizeof(double) * N);
        else {
            memcpy(pop, c->opt, sizeof(double) * N);
            eq_error = best_eq_error;
        }
        iter_nb++;
        found = rel_error < min_error;
    } while (iter_nb < max_iter && !found);
    if (st) {
        st->iters = iter_nb; st->converged = found;
        st->eq_error = eq_error; st->pop_error = pop_error; st->rel_error = rel_error;
    }
    return found;
}

/* ------------------------------------------------------------------------------------------------------- */
/* exported entry points (ctypes); same shapes as oracle/ref_harness.cpp                                    */
/* ------------------------------------------------------------------------------------------------------- */
int ora_abi_version(void) { return 1; }

int ora_lu_solve(double* a, double* b, int n) { lu_solve(a, b, n); return 0; }

int ora_esc_func(const rt_model_desc* d, int n, const double* gamma, const double* delta, double* out)
{
    for (int i = 0; i < n; i++) out[i] = esc_func(&d->lvg, gamma[i], delta[i]);
    return 0;
}

int ora_overlap_esc_func(const rt_model_desc* d, int which, int n, const double* gamma, const double* delta,
                         const double* gratio, const double* dx, double* out)
{
    if (d->scheme != RT_SCHEME_LINE_OVERLAP) return 1;
    ora_model* m = model_create(d);
    for (int i = 0; i < n; i++) out[i] = overlap_esc_func(&m->ov[which == 2 ? 1 : 0], gamma[i], delta[i], gratio[i], dx[i]);
    model_destroy(m);
    return 0;
}

int ora_dust_absorption(const rt_model_desc* d, const rt_layer_soa* lay, int l, int n, const double* energy, double* out)
{
    ora_model* m = model_create(d);
    for (int i = 0; i < n; i++) out[i] = dust_absorption(m, energy[i], lay->dust_conc + (size_t)l * lay->dust_ncomp);
    model_destroy(m);
    return 0;
}

int ora_line_groups(const rt_model_desc* d, int cap, int* groups, int* n_out)
{
    int* g;
    int n = build_line_groups(d, &g);
    *n_out = n;
    memcpy(groups, g, sizeof(int) * 5 * (n < cap ? n : cap));
    free(g);
    return 0;
}

int ora_collision_rates(const rt_model_desc* d, const rt_layer_soa* lay, double* rates)
{
    ora_model* m = model_create(d);
    const int N = m->N;
    for (int l = 0; l < lay->nb_lay; l++) {
        ora_layer s;
        layer_set(m, lay, l, &s);
        double* R = rates + (size_t)l * N * N;
        memset(R, 0, sizeof(double) * (size_t)N * N);
        for (int i = 1; i < N; i++)
            for (int j = 0; j < i; j++) {
                double dn, up, a, b;
                rate_neutrals(m, &s, i, j, &dn, &up);
                rate_electrons(m, &s, i, j, &a, &b);
                R[(size_t)i * N + j] = dn + a;
                R[(size_t)j * N + i] = up + b;
            }
    }
    model_destroy(m);
    return 0;
}

int ora_assemble(const rt_model_desc* d, const rt_layer_soa* lay, const double* pop, double* matrix, double* df)
{
    ora_model* m = model_create(d);
    const int N = m->N;
    double* M = (double*)malloc(sizeof(double) * (size_t)N * N);
    double* f = (double*)malloc(sizeof(double) * N);
    for (int l = 0; l < lay->nb_lay; l++) {
        ora_layer s;
        layer_set(m, lay, l, &s);
        assemble(m, &s, pop + (size_t)l * N, M, f);
        if (matrix) memcpy(matrix + (size_t)l * N * N, M, sizeof(double) * (size_t)N * N);
        if (df) memcpy(df + (size_t)l * N, f, sizeof(double) * N);
    }
    free(M); free(f);
    model_destroy(m);
    return 0;
}

int ora_step(const rt_model_desc* d, const rt_layer_soa* lay, const double* old_pop, double* new_pop, double* eq_err)
{
    ora_model* m = model_create(d);
    const int N = m->N;
    double* work = (double*)malloc(sizeof(double) * ((size_t)N * N + 2 * N));
    for (int l = 0; l < lay->nb_lay; l++) {
        ora_layer s;
        layer_set(m, lay, l, &s);
        calc_new_pop(m, &s, old_pop + (size_t)l * N, new_pop + (size_t)l * N, &eq_err[l], work);
    }
    free(work);
    model_destroy(m);
    return 0;
}

int ora_intensity(const rt_model_desc* d, const rt_layer_soa* lay, int l, int n, const int* up, const int* low,
                  const double* pop, double* out)
{
    ora_model* m = model_create(d);
    ora_layer s;
    layer_set(m, lay, l, &s);
    for (int i = 0; i < n; i++) out[i] = intensity_any(m, &s, up[i], low[i], pop);
    model_destroy(m);
    return 0;
}

int ora_boundary(const rt_model_desc* d, const rt_layer_soa* lay, double* pop)
{
    ora_model* m = model_create(d);
    const int N = m->N;
    double* work = (double*)malloc(sizeof(double) * (size_t)N * N);
    for (int l = 0; l < lay->nb_lay; l++) {
        ora_layer s;
        layer_set(m, lay, l, &s);
        boundary_populations(m, &s, pop + (size_t)l * N, work);
    }
    free(work);
    model_destroy(m);
    return 0;
}

int ora_boltzmann(const rt_model_desc* d, const rt_layer_soa* lay, double* pop)
{
    ora_model* m = model_create(d);
    for (int l = 0; l < lay->nb_lay; l++) boltzmann(m, lay->temp_n[l], pop + (size_t)l * m->N);
    model_destroy(m);
    return 0;
}

static int is_methanol(const rt_model_desc* d)
{
    return d->mol_name && (strcmp(d->mol_name, "CH3OHa") == 0 || strcmp(d->mol_name, "CH3OHe") == 0);
}

/* Independent-start solve of layers [lay_begin, lay_end) on nthreads OpenMP threads (layers are independent);
 * start_mode RT_START_CHAINED runs the serial warm-start chain of calc_molecular_populations
 * (radiative_transfer.cpp:236-281) over ALL layers, retry policy :259-276 included when retry != 0. */
int ora_solve(const rt_model_desc* d, const rt_layer_soa* lay, double* pop, int start_mode, int max_iter, double min_err,
              int accel, int retry, int max_iter_ext, rt_layer_status* st, int lay_begin, int lay_end, int nthreads,
              double* seconds)
{
    ora_model* m = model_create(d);
    const int N = m->N;
    if (lay_end < 0 || lay_end > lay->nb_lay) lay_end = lay->nb_lay;
    if (nthreads < 1) nthreads = 1;
    const int chained = start_mode == RT_START_CHAINED;
    if (chained) { nthreads = 1; lay_begin = 0; lay_end = lay->nb_lay; }
    const double t0 = omp_get_wtime();
#pragma omp parallel num_threads(nthreads)
    {
        double* work = (double*)malloc(sizeof(double) * ((size_t)N * N + 3 * N));
        double* given = (double*)malloc(sizeof(double) * N);
        ora_ctl c;
        ctl_init(&c, N);
        int prev_found = 0;
#pragma omp for schedule(dynamic, 1)
        for (int l = lay_begin; l < lay_end; l++) {
            ora_layer s;
            layer_set(m, lay, l, &s);
            double* p = pop + (size_t)l * N;
            memcpy(given, p, sizeof(double) * N);
            rt_layer_status ls;
            for (int attempt = 0; attempt < 2; attempt++) {
                if (chained) {
                    if (l > 0 && prev_found) memcpy(p, p - N, sizeof(double) * N);
                    else boundary_populations(m, &s, p, work);
                }
                else if (start_mode == RT_START_BOUNDARY) boundary_populations(m, &s, p, work);
                else if (start_mode == RT_START_LTE) boltzmann(m, s.temp_n, p);
                else memcpy(p, given, sizeof(double) * N);
                const int acc = attempt == 0 ? accel : 0;
                const int found = calculate_populations(m, &s, p, attempt == 0 ? max_iter : max_iter_ext, min_err, acc, &ls, work, &c);
                if (found || !(accel && retry && !is_methanol(d)) || attempt == 1) break;
            }
            prev_found = ls.converged;
            if (st) st[l] = ls;
        }
        free(work); free(given); free(c.buf);
    }
    if (seconds) *seconds = omp_get_wtime() - t0;
    model_destroy(m);
    return 0;
}

/* The line whose gain uses the widened thermal width of the oH2O 6_16 -> 5_23 (22.2 GHz) special case of calc_gain
 * (transition_data.cpp:249-260): vel = sqrt((sqrt(2kT/m) + offset)^2 + v_turb^2), offset = 5e4 cm/s in the reference.
 * The reference recognises the line by its quantum numbers (j, k1, k2), which the flat model does not carry, so the
 * caller names it by level indices.  up < 0: no such line. */
static int g_gw_up = -1, g_gw_low = -1;
static double g_gw_offset = 0.;
void ora_set_gain_width_line(int up, int low, double thermal_offset) { g_gw_up = up; g_gw_low = low; g_gw_offset = thermal_offset; }

/* transition_data_container::calc_inv (transition_data.cpp:210-223), calc_gain (:240-294, incl. the special line
 * width above) and calc_line_profile (:296-378) for listed transitions.  scalars [nb_trans][5] = {inv, gain, tau_eff, tau_max,
 * lay_nb_hg}.  nb_freq / nb_aspect / delta_aspect / velocity_shift: 300 / 37 / 0.25 / 5e5 in the reference
 * (transition_data.cpp:16-18,191). */
int ora_maser_postpass(const rt_model_desc* d, const rt_layer_soa* lay, const double* pop, int nb_trans, const int* up,
                       const int* low, int nb_freq, int nb_aspect, double delta_aspect, double velocity_shift,
                       double* inv_arr, double* gain_arr, double* scalars, double* tau_vs_aspect, double* tau_vs_freq)
{
    ora_model* m = model_create(d);
    const int N = m->N, L = lay->nb_lay;
    const rt_params* pr = &d->params;
    double* inv_l = (double*)malloc(sizeof(double) * L);
    double* gain_l = (double*)malloc(sizeof(double) * L);
    double* line_opacity = (double*)malloc(sizeof(double) * L);
    double* dust_opacity = (double*)malloc(sizeof(double) * L);
    double* vel_width = (double*)malloc(sizeof(double) * L);
    double* depth = (double*)malloc(sizeof(double) * (size_t)nb_freq * nb_aspect);
    for (int t = 0; t < nb_trans; t++) {
        const int u = up[t], l = low[t];
        const double gu = d->g[u], gl = d->g[l];
        const double A = d->einst[(size_t)u * N + l];
        const double energy = d->energy[u] - d->energy[l]; /* transition::energy, spectroscopy.cpp:1224-1238 */
        const double energy_th = energy * energy * energy;
        double inv = 0., gain = 0., tau_eff = 0., g = 0.;
        int hg = 0;
        for (int k = 0; k < L; k++) { /* calc_inv */
            inv_l[k] = pop[(size_t)k * N + u] / gu - pop[(size_t)k * N + l] / gl;
            inv += inv_l[k] * lay->dz[k];
        }
        inv /= lay->height;
        for (int k = 0; k < L; k++) { /* calc_gain :248-272 */
            double vel;
            if (u == g_gw_up && l == g_gw_low) /* :257-260 */
                vel = sqrt(pow(sqrt(2. * pr->boltzmann_constant * lay->temp_n[k] / d->mol_mass) + g_gw_offset, 2.)
                           + lay->vel_turb[k] * lay->vel_turb[k]);
            else
                vel = pow(2. * pr->boltzmann_constant * lay->temp_n[k] / d->mol_mass + lay->vel_turb[k] * lay->vel_turb[k], 0.5);
            const double line_gain = inv_l[k] * gu * A * pr->onedivby_sqrt_pi * lay->mol_conc[k] / (energy_th * pr->eight_pi * vel);
            const double d_abs = dust_absorption(m, energy, lay->dust_conc + (size_t)k * lay->dust_ncomp);
            gain_l[k] = line_gain - d_abs;
        }
        for (int k = 0; k < L; k++) { /* :274-289 */
            gain += gain_l[k] * lay->dz[k];
            if (gain_l[k] > 0.) tau_eff += gain_l[k] * lay->dz[k];
            if (gain_l[k] > g) { g = gain_l[k]; hg = k; }
        }
        gain /= lay->height;
        if (g < 1.e-99) hg = 0;
        /* calc_line_profile */
        const double vmax = lay->vel_n[0] + velocity_shift, vmin = lay->vel_n[L - 1] - velocity_shift;
        const double dv = (vmax - vmin) / (nb_freq - 1.);
        for (int k = 0; k < L; k++) {
            vel_width[k] = pow(2. * pr->boltzmann_constant * lay->temp_n[k] / d->mol_mass + lay->vel_turb[k] * lay->vel_turb[k], 0.5);
            line_opacity[k] = inv_l[k] * gu * A * lay->mol_conc[k] * pr->onedivby_sqrt_pi / (energy_th * pr->eight_pi * vel_width[k]);
            dust_opacity[k] = dust_absorption(m, energy, lay->dust_conc + (size_t)k * lay->dust_ncomp);
        }
        for (int i = 0; i < nb_aspect; i++) {
            const double aspect = 1. + delta_aspect * i;
            double vel = vmin;
            for (int n = 0; n < nb_freq; vel += dv, n++) {
                double tau = 0.;
                for (int k = 0; k < L; k++) {
                    const double x = (vel - lay->vel_n[k] / aspect) / vel_width[k];
                    const double profile = exp(-x * x);
                    if (line_opacity[k] * profile - dust_opacity[k] > 0.)
                        tau += (line_opacity[k] * profile - dust_opacity[k]) * lay->dz[k] * aspect;
                }
                depth[(size_t)n * nb_aspect + i] = tau;
            }
        }
        for (int i = 0; i < nb_aspect; i++) {
            double x = 0.;
            for (int n = 0; n < nb_freq; n++)
                if (x < depth[(size_t)n * nb_aspect + i]) x = depth[(size_t)n * nb_aspect + i];
            if (tau_vs_aspect) tau_vs_aspect[(size_t)t * nb_aspect + i] = x;
            if (i == 0 && scalars) scalars[5 * (size_t)t + 3] = x;
        }
        if (tau_vs_freq)
            for (int n = 0; n < nb_freq; n++) tau_vs_freq[(size_t)t * nb_freq + n] = depth[(size_t)n * nb_aspect];
        if (inv_arr) memcpy(inv_arr + (size_t)t * L, inv_l, sizeof(double) * L);
        if (gain_arr) memcpy(gain_arr + (size_t)t * L, gain_l, sizeof(double) * L);
        if (scalars) { double* s = scalars + 5 * (size_t)t; s[0] = inv; s[1] = gain; s[2] = tau_eff; s[4] = hg; }
    }
    free(inv_l); free(gain_l); free(line_opacity); free(dust_opacity); free(vel_width); free(depth);
    model_destroy(m);
    return 0;
}

/* lim_luminosity_lvg (maser_luminosity.cpp:7-106), calc_exc_temp (transition_data.cpp:225-237) and
 * calc_saturation_depth (transition_data.cpp:456-483) for listed transitions.  Per-layer outputs [nb_trans][L].
 * layer0_intensity != 0: the reference's literal behaviour -- intensity_calc receives the base pointer of the population
 * array (maser_luminosity.cpp:54,58), i.e. J of the loss terms is evaluated with the populations of layer 0.
 * lay_nb_hg [nb_trans]: input (layer of the highest gain from ora_maser_postpass), NULL skips the saturation part. */
int ora_lim_luminosity(const rt_model_desc* d, const rt_layer_soa* lay, const double* pop, int nb_trans, const int* up, const int* low,
                       int layer0_intensity, double* lum_arr, double* pump_eff_arr, double* loss_rate_arr, double* pump_rate_arr,
                       double* emiss_coeff_arr, double* exc_temp_arr, double* lum, const int* lay_nb_hg, double beaming_factor,
                       double* tau_sat, double* temp_sat)
{
    ora_model* m = model_create(d);
    const int N = m->N, L = lay->nb_lay;
    const rt_params* pr = &d->params;
    double* up_loss = (double*)malloc(sizeof(double) * L);
    double* low_loss = (double*)malloc(sizeof(double) * L);
    for (int t = 0; t < nb_trans; t++) {
        const int lo = low[t], hi = up[t];
        const double gu = d->g[hi], gl = d->g[lo];
        const double energy = d->energy[hi] - d->energy[lo];
        double lum_t = 0.;
        for (int l = 0; l < L; l++) {
            const double* pl = pop + (size_t)l * N;
            const double* pj = layer0_intensity ? pop : pl;
            ora_layer s;
            layer_set(m, lay, l, &s);
            for (int k = 0; k <= 1; k++) {
                const int j = k == 0 ? lo : hi;
                double loss = 0.;
                for (int i = 0; i < N; i++) {
                    const double Aij = d->einst[(size_t)i * N + j], Aji = d->einst[(size_t)j * N + i];
                    if (Aij != 0. && i != lo && i != hi) {
                        if (i < j && pl[i] * Aij > pl[j] * Aji) loss += Aji * (1. + intensity_any(m, &s, j, i, pj));
                        else if (i > j && pl[i] * Aij < pl[j] * Aji) loss += Aji * intensity_any(m, &s, i, j, pj);
                    }
                }
                if (k == 0) low_loss[l] = loss;
                else up_loss[l] = loss;
            }
            for (int i = 0; i < N; i++) { /* get_rate_neutrals(init, fin), coll_rates.cpp:225-240 */
                double dn, ur;
                if (i != lo) {
                    if (lo > i) { rate_neutrals(m, &s, lo, i, &dn, &ur); low_loss[l] += dn; }
                    else { rate_neutrals(m, &s, i, lo, &dn, &ur); low_loss[l] += ur; }
                }
                if (i != hi) {
                    if (hi > i) { rate_neutrals(m, &s, hi, i, &dn, &ur); up_loss[l] += dn; }
                    else { rate_neutrals(m, &s, i, hi, &dn, &ur); up_loss[l] += ur; }
                }
            }
            const size_t o = (size_t)t * L + l;
            const double h2 = lay->ph2_conc[l] + lay->oh2_conc[l];
            const double inversion = pl[hi] / gu - pl[lo] / gl;
            double lum_l, eff_l;
            if (inversion > 0.) {
                lum_l = inversion / (1. / (up_loss[l] * gu) + 1. / (low_loss[l] * gl)) * lay->mol_conc[l];
                eff_l = inversion / (pl[hi] / gu + pl[lo] / gl);
            }
            else lum_l = eff_l = 1.e-99;
            lum_t += lum_l * lay->dz[l];
            if (lum_arr) lum_arr[o] = lum_l;
            if (pump_eff_arr) pump_eff_arr[o] = eff_l;
            if (emiss_coeff_arr) emiss_coeff_arr[o] = h2 * lay->mol_conc[l] / lay->velg_n[l];
            if (loss_rate_arr) loss_rate_arr[o] = (up_loss[l] * gu + low_loss[l] * gl) / (gu + gl);
            if (pump_rate_arr) pump_rate_arr[o] = 0.5 * (pl[hi] * up_loss[l] + pl[lo] * low_loss[l]) / h2;
            if (exc_temp_arr) exc_temp_arr[o] = pr->cm_inverse_to_kelvins * energy / log((pl[lo] * gu) / (pl[hi] * gl));
        }
        if (lum) lum[t] = lum_t / lay->height;
        if (lay_nb_hg && (tau_sat || temp_sat)) {
            const int l = lay_nb_hg[t];
            const double* pl = pop + (size_t)l * N;
            const double exc_temp = pr->cm_inverse_to_kelvins * energy / log((pl[lo] * gu) / (pl[hi] * gl));
            double ts = 0., tmp = 0.;
            if (exc_temp < 0.) {
                const double source_function = 1. / (1. - exp(energy * pr->cm_inverse_to_kelvins / exc_temp));
                const double cmb_intensity = 1. / (exp(energy * pr->cm_inverse_to_kelvins / 2.73) - 1.);
                const double intensity_0 = source_function + cmb_intensity;
                const double loss = (up_loss[l] * gu + low_loss[l] * gl) / (gu + gl);
                ts = log(loss / (d->einst[(size_t)hi * N + lo] * beaming_factor * intensity_0));
                tmp = intensity_0 * exp(ts) * pr->planck_constant * (pr->speed_of_light * energy) / pr->boltzmann_constant;
            }
            if (tau_sat) tau_sat[t] = ts;
            if (temp_sat) temp_sat[t] = tmp;
        }
    }
    free(up_loss); free(low_loss);
    model_destroy(m);
    return 0;
}
