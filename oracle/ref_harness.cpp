// oracle/ref_harness.cpp -- TEST INFRASTRUCTURE ONLY (never linked into librt_b200.so).
//
// Drives the UNMODIFIED reference sources (compiled where they lie under /root/reference by oracle/Makefile,
// against the restated math headers in oracle/shim/) from the same flat problem description the CUDA library
// takes (include/rt_b200.h).  Output: oracle/_ref/libref_oracle.so, loaded with ctypes by tests/ and by
// bench.py's cpu_baseline / --impl reference legs.
//
// What is reference code and what is harness:
//   * reference (untouched): iteration_scheme_lvg / iteration_scheme_line_overlap, iteration_control<T>,
//     boundary_layer_populations, calc_molecular_populations, lvg_method_data(+overlap), dust_component /
//     dust_model::absorption, collision_data::get_rate/locate, the species mixing rules
//     ch3oh_collisions / oh_hf_collisions / h2o_collisions ::get_rate_neutrals + set_gas_param,
//     transition_data_container::calc_inv/calc_gain/calc_line_profile.
//   * harness (this file): synthetic subclasses that FILL the reference's containers from flat arrays instead
//     of parsing the (absent) data files, and temp text files in the reference's own formats for the loaders
//     that only read files (lvg tables: lvg_method_functions.cpp:37-56,281-308; dust: dust_model.cpp:159-185).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <string>
#include <vector>
#include <sstream>
#include <fstream>
#include <iostream>
#include <unistd.h>
#include <sys/stat.h>
#include <omp.h>

#include "constants.h"
#include "utils.h"
#include "linear_algebra.h"
#include "spectroscopy.h"
#include "coll_rates.h"
#include "coll_rates_ch3oh.h"
#include "coll_rates_oh.h"
#include "coll_rates_h2o.h"
#include "coll_rates_nh3.h"
#include "coll_rates_h2co.h"
#include "lvg_method_functions.h"
#include "dust_model.h"
#include "cloud_data.h"
#include "iteration_lvg.h"
#include "iteration_control.h"
#include "transition_data.h"
#include "maser_luminosity.h"
#include <list>

#include "../include/rt_b200.h"

// linked from the reference's radiative_transfer.cpp (compiled with -Dmain=reference_main)
void calc_molecular_populations(cloud_data* cloud, iteration_scheme_lvg* it_scheme_lvg, energy_diagram* mol_levels,
                                einstein_coeff* mol_einst, collisional_transitions* mol_coll, double* mol_popul,
                                int nb_lev, bool acceleration, int verbosity);

// The flat model carries no k1 / k2 quantum numbers.  To exercise the oH2O 6_16 -> 5_23 special line width of
// calc_gain (transition_data.cpp:249-260) the caller names one line by level indices; those two levels then get the
// quantum numbers the reference tests for (v = 0, j/k1/k2 = 6/1/6 and 5/2/3).  up < 0: none.
static int g_gw_up = -1, g_gw_low = -1;
extern "C" void ref_set_gain_width_line(int up, int low) { g_gw_up = up; g_gw_low = low; }

namespace {

// ---------------------------------------------------------------------------------------------------------
// synthetic containers
// ---------------------------------------------------------------------------------------------------------
struct synth_diagram : public energy_diagram {
    synth_diagram(const rt_model_desc* d)
        : energy_diagram(molecule(d->mol_name ? d->mol_name : "synthetic", 1, d->mol_mass, 0.), 0)
    {
        for (int i = 0; i < d->nb_lev; i++) {
            energy_level lev;
            lev.nb = i;
            lev.g = d->g[i];
            lev.v = d->v ? d->v[i] : 0;
            lev.j = d->j ? d->j[i] : 0.;
            lev.energy = d->energy[i];
            if (i == g_gw_up) { lev.v = 0; lev.j = 6; lev.k1 = 1; lev.k2 = 6; }
            if (i == g_gw_low) { lev.v = 0; lev.j = 5; lev.k1 = 2; lev.k2 = 3; }
            lev_array.push_back(lev);
        }
        nb_lev = d->nb_lev;
    }
};

struct synth_einstein : public einstein_coeff {
    synth_einstein(const energy_diagram* di, const rt_model_desc* d) : einstein_coeff(di)
    {
        memcpy(*arr, d->einst, sizeof(double) * (size_t)nb_lev * nb_lev);
    }
};

// fills the members of either table class of the reference and runs ITS calc_coeff_deriv (virtual)
template <class Base> struct synth_coll_data_t : public Base {
    using Base::nb_lev; using Base::jmax; using Base::imax; using Base::tgrid; using Base::coeff; using Base::coeff_deriv;
    synth_coll_data_t(const rt_coll_table* t)
    {
        nb_lev = t->nb_lev;
        jmax = t->jmax;
        imax = nb_lev * (nb_lev - 1) / 2;
        tgrid = new double[jmax];
        memcpy(tgrid, t->tgrid, sizeof(double) * jmax);
        coeff = alloc_2d_array<double>(imax, jmax);
        memcpy(*coeff, t->coeff, sizeof(double) * (size_t)imax * jmax);
        coeff_deriv = alloc_2d_array<double>(imax, jmax);
        memset(*coeff_deriv, 0, sizeof(double) * (size_t)imax * jmax);
        Base::calc_coeff_deriv(); // the reference's own rule: linear coll_rates.cpp:61-69, cubic spline :89-119
    }
};
typedef synth_coll_data_t<collision_data> synth_coll_data;
typedef synth_coll_data_t<collision_data_cub_spline> synth_coll_data_spline;
static collision_data* make_coll_data(const rt_coll_table* t)
{
    if (t->interp == RT_INTERP_CUBIC_SPLINE) return new synth_coll_data_spline(t);
    return new synth_coll_data(t);
}

// A collisional_transitions whose tables are synthetic but whose MIXING RULES are the reference's own compiled
// member functions.  ch3oh_collisions / oh_hf_collisions / h2o_collisions add no data members and their
// constructors only parse files, so their rule methods are invoked non-virtually (qualified call) on this
// object: they touch nothing but the base-class members coll_data / max_temp / nb1..nb3 filled below.
struct synth_collisions : public collisional_transitions {
    int rule;
    synth_collisions(const rt_model_desc* d) : rule(d->species_rule)
    {
        nb_lev = d->nb_lev;
        nb1 = d->nb_neutral_tables;
        nb2 = nb1 + d->nb_electron_tables;
        nb3 = nb2;
        for (int k = 0; k < nb3; k++) coll_data.push_back(make_coll_data(&d->tables[k]));
        max_temp = new double[nb3 > 0 ? nb3 : 1];
        for (int k = 0; k < nb3; k++) max_temp[k] = coll_data[k]->get_max_temp();
    }
    void set_gas_param(double tn, double te, double he, double ph2, double oh2, double h, double el,
                       double*& conc, int*& ind) const override
    {
        switch (rule) {
        case RT_RULE_CH3OH:
            reinterpret_cast<const ch3oh_collisions*>(this)->ch3oh_collisions::set_gas_param(tn, te, he, ph2, oh2, h, el, conc, ind);
            break;
        case RT_RULE_OH_HF:
            reinterpret_cast<const oh_hf_collisions*>(this)->oh_hf_collisions::set_gas_param(tn, te, he, ph2, oh2, h, el, conc, ind);
            break;
        case RT_RULE_H2O:
            reinterpret_cast<const h2o_collisions*>(this)->h2o_collisions::set_gas_param(tn, te, he, ph2, oh2, h, el, conc, ind);
            break;
        case RT_RULE_NH3:
            reinterpret_cast<const nh3_collisions*>(this)->nh3_collisions::set_gas_param(tn, te, he, ph2, oh2, h, el, conc, ind);
            break;
        case RT_RULE_H2CO:
            reinterpret_cast<const h2co_collisions*>(this)->h2co_collisions::set_gas_param(tn, te, he, ph2, oh2, h, el, conc, ind);
            break;
        default:
            // generic: base class fills indices (+ electron conc); neutral partner k gets, in order,
            // He, p-H2, o-H2, H (the convention every species class of the reference follows for its first tables)
            collisional_transitions::set_gas_param(tn, te, he, ph2, oh2, h, el, conc, ind);
            {
                const double c[4] = {he, ph2, oh2, h};
                for (int k = 0; k < nb1; k++) conc[k] = c[k < 4 ? k : 3];
            }
        }
    }
    void get_rate_neutrals(const energy_level& up, const energy_level& low, double& down, double& upr, double tn,
                           const double* conc, const int* ind) const override
    {
        switch (rule) {
        case RT_RULE_CH3OH:
            reinterpret_cast<const ch3oh_collisions*>(this)->ch3oh_collisions::get_rate_neutrals(up, low, down, upr, tn, conc, ind);
            break;
        case RT_RULE_OH_HF:
            reinterpret_cast<const oh_hf_collisions*>(this)->oh_hf_collisions::get_rate_neutrals(up, low, down, upr, tn, conc, ind);
            break;
        case RT_RULE_H2O:
            reinterpret_cast<const h2o_collisions*>(this)->h2o_collisions::get_rate_neutrals(up, low, down, upr, tn, conc, ind);
            break;
        case RT_RULE_NH3:
            reinterpret_cast<const nh3_collisions*>(this)->nh3_collisions::get_rate_neutrals(up, low, down, upr, tn, conc, ind);
            break;
        case RT_RULE_H2CO:
            reinterpret_cast<const h2co_collisions*>(this)->h2co_collisions::get_rate_neutrals(up, low, down, upr, tn, conc, ind);
            break;
        default:
            collisional_transitions::get_rate_neutrals(up, low, down, upr, tn, conc, ind);
        }
    }
};

// ---------------------------------------------------------------------------------------------------------
// temp files in the reference's own text formats (17 significant digits: exact round trip through strtod)
// ---------------------------------------------------------------------------------------------------------
std::string make_tmpdir()
{
    const char* base = getenv("RT_ORACLE_TMP");
    std::string tmpl = std::string(base ? base : "/tmp") + "/rt_ref_XXXXXX";
    std::vector<char> buf(tmpl.begin(), tmpl.end());
    buf.push_back(0);
    if (!mkdtemp(buf.data())) { perror("mkdtemp"); return ""; }
    std::string d(buf.data());
    mkdir((d + "/lvg").c_str(), 0700);
    return d + "/";
}

void rm_tmpdir(const std::string& d)
{
    if (d.empty()) return;
    const char* f[] = {"lvg/lvg_loss_func.txt", "lvg/line_overlap_func_p1.txt", "lvg/line_overlap_func_p2.txt"};
    for (auto s : f) unlink((d + s).c_str());
    for (int i = 0; i < 16; i++) unlink((d + "dust_" + std::to_string(i) + ".txt").c_str());
    rmdir((d + "lvg").c_str());
    rmdir(d.c_str());
}

// lvg_method_functions.cpp:37-56
void write_lvg_table(const std::string& fname, const rt_lvg_table& t)
{
    FILE* f = fopen(fname.c_str(), "w");
    fprintf(f, "! synthetic LVG loss function\n! delta rows, gamma columns\n! nb_d nb_g\n%d %d\n", t.nb_d, t.nb_g);
    for (int i = 0; i < t.nb_g; i++) fprintf(f, "%.17g ", t.gamma[i]);
    fprintf(f, "\n");
    for (int i = 0; i < t.nb_d; i++) {
        fprintf(f, "%d %.17g", i, t.delta[i]);
        for (int j = 0; j < t.nb_g; j++) fprintf(f, " %.17g", t.p[(size_t)i * t.nb_g + j]);
        fprintf(f, "\n");
    }
    fclose(f);
}

// lvg_method_functions.cpp:281-308
void write_overlap_table(const std::string& fname, const rt_overlap_table& t)
{
    FILE* f = fopen(fname.c_str(), "w");
    fprintf(f, "! synthetic line overlap function\n!\n! nb_d nb_dx nb_gr nb_g\n%d %d %d %d\n", t.nb_d, t.nb_dx, t.nb_gr, t.nb_g);
    for (int i = 0; i < t.nb_g; i++) fprintf(f, "%.17g ", t.gamma[i]);
    fprintf(f, "\n");
    for (int l = 0; l < t.nb_d; l++)
        for (int k = 0; k < t.nb_dx; k++) {
            fprintf(f, "%.17g %.17g\n", t.delta[l], t.dx[k]);
            for (int i = 0; i < t.nb_gr; i++) {
                fprintf(f, "%d %.17g", i, t.gratio[i]);
                const double* row = t.p + (size_t)(l * t.nb_dx + k) * ((size_t)t.nb_gr * t.nb_g) + (size_t)i * t.nb_g;
                for (int j = 0; j < t.nb_g; j++) fprintf(f, " %.17g", row[j]);
                fprintf(f, "\n");
            }
        }
    fclose(f);
}

// dust_model.cpp:159-185.  The loader computes ph_en = 1e4/lambda and abs_coeff = k_abs*dust_mass_perH/conc_perH;
// with dust_mass_perH = conc_perH = 1 the second is exact, and lambda is searched so that 1e4/lambda == E bit for bit.
bool write_dust_file(const std::string& fname, const rt_dust_comp& c)
{
    FILE* f = fopen(fname.c_str(), "w");
    for (int i = 0; i < 6; i++) fprintf(f, "! synthetic dust component, comment line %d\n", i);
    fprintf(f, "synth 1 1 1 1 %.17g x\n", c.wvl_exp);
    fprintf(f, "%d\n", c.nb_ph_en);
    bool exact = true;
    for (int i = 0; i < c.nb_ph_en; i++) {
        const double e = c.ph_en[i];
        double lam = 1.e+4 / e;
        if (1.e+4 / lam != e) {
            bool ok = false;
            double lo = lam, hi = lam;
            for (int s = 0; s < 8 && !ok; s++) {
                lo = nextafter(lo, 0.);
                hi = nextafter(hi, 1e300);
                if (1.e+4 / lo == e) { lam = lo; ok = true; }
                else if (1.e+4 / hi == e) { lam = hi; ok = true; }
            }
            if (!ok) exact = false;
        }
        fprintf(f, "%.17g 0 0 0 %.17g 0\n", lam, c.abs_coeff[i]);
    }
    fclose(f);
    return exact;
}

struct null_buf : public std::streambuf {
    int overflow(int c) override { return c; }
};

// RAII: silence / capture the reference's cout chatter
struct cout_redirect {
    std::streambuf* old;
    cout_redirect(std::streambuf* b) { old = std::cout.rdbuf(b); }
    ~cout_redirect() { std::cout.rdbuf(old); }
};

// Forwarding scheme for iteration_control<T>'s duck type (get_vector_dim / calc_new_pop): counts steps.
struct counting_scheme {
    iteration_scheme_lvg* s;
    int calls;
    int get_vector_dim() { return s->get_vector_dim(); }
    void calc_new_pop(double* o, double* n, double& e) { calls++; s->calc_new_pop(o, n, e); }
};

// exposes the protected matrix of iteration_scheme_lvg for step-parity tests
struct lvg_probe : public iteration_scheme_lvg {
    using iteration_scheme_lvg::iteration_scheme_lvg;
    double** mat() { return matrix; }
};
struct overlap_probe : public iteration_scheme_line_overlap {
    using iteration_scheme_line_overlap::iteration_scheme_line_overlap;
    double** mat() { return matrix; }
    const std::vector<hfs_lines>& groups() const { return line_list; }
};

struct ref_ctx {
    std::string tmp;
    synth_diagram* diagram = 0;
    synth_einstein* einst = 0;
    synth_collisions* coll = 0;
    dust_model* dust = 0;
    cloud_data* cloud = 0;
    iteration_scheme_lvg* scheme = 0; // lvg_probe or overlap_probe
    bool overlap = false;
    bool dust_exact = true;
    int N = 0, L = 0;

    double** mat() { return overlap ? static_cast<overlap_probe*>(scheme)->mat() : static_cast<lvg_probe*>(scheme)->mat(); }

    ~ref_ctx()
    {
        delete scheme;
        delete cloud;
        delete dust;
        delete coll;
        delete einst;
        delete diagram;
        rm_tmpdir(tmp);
    }
    void set_layer(int lay)
    {
        // == radiative_transfer.cpp:238-242
        cloud_layer cl = cloud->lay_array[lay];
        scheme->set_vel_grad(cl.velg_n);
        scheme->set_dust_parameters(cl.dust_grain_conc, cl.dust_grain_temp);
        scheme->set_parameters(cl.temp_n, cl.temp_el, cl.el_conc, cl.h_conc, cl.ph2_conc, cl.oh2_conc, cl.he_conc, cl.mol_conc, cl.vel_turb);
    }
};

ref_ctx* build_ctx(const rt_model_desc* d, const rt_layer_soa* lay)
{
    null_buf nb;
    cout_redirect quiet(&nb);
    ref_ctx* c = new ref_ctx;
    c->N = d->nb_lev;
    c->tmp = make_tmpdir();
    c->diagram = new synth_diagram(d);
    c->einst = new synth_einstein(c->diagram, d);
    c->coll = new synth_collisions(d);

    c->dust = new dust_model(0);
    for (int i = 0; i < d->dust_ncomp; i++) {
        std::string fn = c->tmp + "dust_" + std::to_string(i) + ".txt";
        c->dust_exact = write_dust_file(fn, d->dust[i]) && c->dust_exact;
        c->dust->add_component(new dust_component(fn, 0));
    }

    write_lvg_table(c->tmp + "lvg/lvg_loss_func.txt", d->lvg);
    c->overlap = (d->scheme == RT_SCHEME_LINE_OVERLAP);
    if (c->overlap) {
        write_overlap_table(c->tmp + "lvg/line_overlap_func_p1.txt", *d->overlap1);
        write_overlap_table(c->tmp + "lvg/line_overlap_func_p2.txt", *d->overlap2);
        c->scheme = new overlap_probe(c->tmp, 0);
    }
    else c->scheme = new lvg_probe(c->tmp, 0);
    c->scheme->set_dust_model(c->dust);
    c->scheme->init_molecule_data(c->diagram, c->einst, c->coll);

    c->cloud = new cloud_data;
    if (lay) {
        c->L = lay->nb_lay;
        double z = 0.;
        for (int l = 0; l < lay->nb_lay; l++) {
            cloud_layer cl;
            cl.temp_n = lay->temp_n[l]; cl.temp_el = lay->temp_el[l];
            cl.el_conc = lay->el_conc[l]; cl.h_conc = lay->h_conc[l];
            cl.ph2_conc = lay->ph2_conc[l]; cl.oh2_conc = lay->oh2_conc[l];
            cl.he_conc = lay->he_conc[l]; cl.mol_conc = lay->mol_conc[l];
            cl.vel_turb = lay->vel_turb[l]; cl.velg_n = lay->velg_n[l];
            cl.dz = lay->dz ? lay->dz[l] : 0.; cl.vel_n = lay->vel_n ? lay->vel_n[l] : 0.;
            cl.zl = z; z += cl.dz; cl.zu = z; cl.zm = 0.5 * (cl.zl + cl.zu);
            for (int k = 0; k < lay->dust_ncomp; k++) {
                cl.dust_grain_conc.push_back(lay->dust_conc[(size_t)l * lay->dust_ncomp + k]);
                cl.dust_grain_temp.push_back(20.);
            }
            c->cloud->add_layer(cl);
        }
        // get_height() = last.zu - first.zl: make it equal the caller's number bit for bit
        if (lay->nb_lay > 0 && lay->dz) c->cloud->lay_array.back().zu = c->cloud->lay_array.front().zl + lay->height;
        c->cloud->set_dust_model(c->dust);
    }
    return c;
}

} // namespace

// =========================================================================================================
// C entry points (ctypes).  All return 0 on success.
// =========================================================================================================
extern "C" {

int ref_abi_version(void) { return 1; }
// rounding-sensitivity probe of the restated lu_matrix_solve (oracle/shim/linear_algebra.h); 0 = the textbook solve
void ref_set_lu_variant(int v) { shim_lu_variant() = v; }

// the restated dense solver itself (oracle/shim/linear_algebra.h), for LU unit tests: a is [n][n] row-major
int ref_lu_solve(double* a, double* b, int n)
{
    double** rows = new double*[n];
    for (int i = 0; i < n; i++) rows[i] = a + (size_t)i * n;
    lu_matrix_solve(rows, b, n);
    delete[] rows;
    return 0;
}

// lvg_method_data::get_esc_func on arrays of (gamma, delta)
int ref_esc_func(const rt_model_desc* d, int n, const double* gamma, const double* delta, double* out)
{
    ref_ctx* c = build_ctx(d, 0);
    null_buf nb; cout_redirect quiet(&nb);
    lvg_method_data t(c->tmp, "lvg/lvg_loss_func.txt", 0);
    for (int i = 0; i < n; i++) out[i] = t.get_esc_func(gamma[i], delta[i]);
    delete c;
    return 0;
}

// lvg_line_overlap_data::get_esc_func (table 1 or 2) on arrays
int ref_overlap_esc_func(const rt_model_desc* d, int which, int n, const double* gamma, const double* delta,
                         const double* gratio, const double* dx, double* out)
{
    if (d->scheme != RT_SCHEME_LINE_OVERLAP) return 1;
    ref_ctx* c = build_ctx(d, 0);
    null_buf nb; cout_redirect quiet(&nb);
    lvg_line_overlap_data t(c->tmp, which == 2 ? "lvg/line_overlap_func_p2.txt" : "lvg/line_overlap_func_p1.txt", 0);
    for (int i = 0; i < n; i++) out[i] = t.get_esc_func(gamma[i], delta[i], gratio[i], dx[i]);
    delete c;
    return 0;
}

// dust_model::absorption(energy, conc) for layer `lay`
int ref_dust_absorption(const rt_model_desc* d, const rt_layer_soa* lay, int l, int n, const double* energy, double* out)
{
    ref_ctx* c = build_ctx(d, lay);
    for (int i = 0; i < n; i++) out[i] = c->dust->absorption(energy[i], c->cloud->lay_array[l].dust_grain_conc);
    int ok = c->dust_exact ? 0 : 2;
    delete c;
    return ok;
}

// line_list of iteration_scheme_line_overlap::init_molecule_data; groups [cap][5] = {nb, u0, l0, u1, l1}
int ref_line_groups(const rt_model_desc* d, int cap, int* groups, int* n_out)
{
    if (d->scheme != RT_SCHEME_LINE_OVERLAP) return 1;
    ref_ctx* c = build_ctx(d, 0);
    const std::vector<hfs_lines>& g = static_cast<overlap_probe*>(c->scheme)->groups();
    *n_out = (int)g.size();
    for (int i = 0; i < (int)g.size() && i < cap; i++) {
        groups[5 * i] = g[i].nb;
        groups[5 * i + 1] = g[i].upl[0]; groups[5 * i + 2] = g[i].lowl[0];
        groups[5 * i + 3] = g[i].nb > 1 ? g[i].upl[1] : -1; groups[5 * i + 4] = g[i].nb > 1 ? g[i].lowl[1] : -1;
    }
    delete c;
    return 0;
}

// collisional matrix part only: down[L][N][N] (entry [i][j], i>j = down rate i->j; [j][i] = up rate j->i),
// neutrals + electrons, exactly the two calls at iteration_lvg.cpp:121-125
int ref_collision_rates(const rt_model_desc* d, const rt_layer_soa* lay, double* rates)
{
    ref_ctx* c = build_ctx(d, lay);
    const int N = c->N;
    for (int l = 0; l < c->L; l++) {
        const cloud_layer& cl = c->cloud->lay_array[l];
        double* conc = 0; int* ind = 0;
        c->coll->set_gas_param(cl.temp_n, cl.temp_el, cl.he_conc, cl.ph2_conc, cl.oh2_conc, cl.h_conc, cl.el_conc, conc, ind);
        double* R = rates + (size_t)l * N * N;
        memset(R, 0, sizeof(double) * N * N);
        for (int i = 1; i < N; i++)
            for (int j = 0; j < i; j++) {
                double dn, up, a, b;
                c->coll->get_rate_neutrals(c->diagram->lev_array[i], c->diagram->lev_array[j], dn, up, cl.temp_n, conc, ind);
                c->coll->get_rate_electrons(c->diagram->lev_array[i], c->diagram->lev_array[j], a, b, cl.temp_el, conc, ind);
                R[(size_t)i * N + j] = dn + a;
                R[(size_t)j * N + i] = up + b;
            }
        delete[] conc; delete[] ind;
    }
    delete c;
    return 0;
}

// iteration_scheme_*::operator() per layer: matrix [L][N][N] (may be 0), df [L][N] (may be 0)
int ref_assemble(const rt_model_desc* d, const rt_layer_soa* lay, const double* pop, double* matrix, double* df)
{
    ref_ctx* c = build_ctx(d, lay);
    const int N = c->N;
    std::vector<double> p(N), f(N);
    for (int l = 0; l < c->L; l++) {
        c->set_layer(l);
        memcpy(p.data(), pop + (size_t)l * N, sizeof(double) * N);
        (*c->scheme)(N, p.data(), f.data());
        if (matrix) memcpy(matrix + (size_t)l * N * N, *c->mat(), sizeof(double) * N * N);
        if (df) memcpy(df + (size_t)l * N, f.data(), sizeof(double) * N);
    }
    delete c;
    return 0;
}

// iteration_scheme_lvg::calc_new_pop per layer
int ref_step(const rt_model_desc* d, const rt_layer_soa* lay, const double* old_pop, double* new_pop, double* eq_err)
{
    ref_ctx* c = build_ctx(d, lay);
    const int N = c->N;
    std::vector<double> p(N);
    for (int l = 0; l < c->L; l++) {
        c->set_layer(l);
        memcpy(p.data(), old_pop + (size_t)l * N, sizeof(double) * N);
        c->scheme->calc_new_pop(p.data(), new_pop + (size_t)l * N, eq_err[l]);
    }
    delete c;
    return 0;
}

// virtual intensity_calc(up, low, pop, J) for a list of lines of one layer
int ref_intensity(const rt_model_desc* d, const rt_layer_soa* lay, int l, int n, const int* up, const int* low,
                  const double* pop, double* out)
{
    ref_ctx* c = build_ctx(d, lay);
    std::vector<double> p(pop, pop + c->N);
    c->set_layer(l);
    for (int i = 0; i < n; i++) c->scheme->intensity_calc(up[i], low[i], p.data(), out[i]);
    delete c;
    return 0;
}

// boundary_layer_populations for every layer
int ref_boundary(const rt_model_desc* d, const rt_layer_soa* lay, double* pop)
{
    ref_ctx* c = build_ctx(d, lay);
    for (int l = 0; l < c->L; l++) {
        const cloud_layer& cl = c->cloud->lay_array[l];
        boundary_layer_populations(pop + (size_t)l * c->N, c->diagram, c->einst, c->coll, cl.temp_n, cl.temp_el, cl.el_conc,
                                   cl.h_conc, cl.ph2_conc, cl.oh2_conc, cl.he_conc);
    }
    delete c;
    return 0;
}

int ref_boltzmann(const rt_model_desc* d, const rt_layer_soa* lay, double* pop)
{
    ref_ctx* c = build_ctx(d, lay);
    for (int l = 0; l < c->L; l++) boltzmann_populations(c->cloud->lay_array[l].temp_n, pop + (size_t)l * c->N, c->diagram);
    delete c;
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Solves.  Independent-start mode: every layer starts from boundary_layer_populations (the reference's own
// path for layer 0 and after a failed layer, radiative_transfer.cpp:250-253), LTE, or a given vector, then
// iteration_control<T>::calculate_populations.  Layers are independent, so `nthreads` OpenMP threads each own a
// private scheme object (the reference's own threading model: private objects per thread,
// radiative_transfer.cpp:152-201); tables are shared read-only.
// retry != 0 applies the policy of radiative_transfer.cpp:259-276.
// ---------------------------------------------------------------------------------------------------------
static iteration_scheme_lvg* make_scheme(ref_ctx* c)
{
    iteration_scheme_lvg* s;
    if (c->overlap) s = new iteration_scheme_line_overlap(c->tmp, 0);
    else s = new iteration_scheme_lvg(c->tmp, 0);
    s->set_dust_model(c->dust);
    s->init_molecule_data(c->diagram, c->einst, c->coll);
    return s;
}

int ref_solve(const rt_model_desc* d, const rt_layer_soa* lay, double* pop, int start_mode, int max_iter, double min_err,
              int accel, int retry, int max_iter_ext, rt_layer_status* st, int lay_begin, int lay_end, int nthreads,
              double* seconds)
{
    if (start_mode == RT_START_CHAINED) return 1;
    ref_ctx* c = build_ctx(d, lay);
    const int N = c->N;
    if (lay_end < 0 || lay_end > c->L) lay_end = c->L;
    if (nthreads < 1) nthreads = 1;
    null_buf nb;
    cout_redirect quiet(&nb);
    const std::string name = c->diagram->mol.name;
    double t0 = 0.;
#pragma omp parallel num_threads(nthreads)
    {
        iteration_scheme_lvg* s;
#pragma omp critical
        s = make_scheme(c);
        counting_scheme cs{s, 0};
        iteration_control<counting_scheme> ctl(&cs);
        // the clock starts when every thread has its private scheme object (the reference builds those once per model,
        // radiative_transfer.cpp:152-201; their serialised construction is set-up, not the solve being timed)
#pragma omp barrier
#pragma omp master
        t0 = omp_get_wtime();
#pragma omp for schedule(dynamic, 1)
        for (int l = lay_begin; l < lay_end; l++) {
            const cloud_layer cl = c->cloud->lay_array[l];
            s->set_vel_grad(cl.velg_n);
            std::vector<double> dc(cl.dust_grain_conc), dt(cl.dust_grain_temp);
            s->set_dust_parameters(dc, dt);
            s->set_parameters(cl.temp_n, cl.temp_el, cl.el_conc, cl.h_conc, cl.ph2_conc, cl.oh2_conc, cl.he_conc, cl.mol_conc, cl.vel_turb);
            double* p = pop + (size_t)l * N;
            std::vector<double> given(p, p + N);
            auto start = [&]() {
                if (start_mode == RT_START_BOUNDARY)
                    boundary_layer_populations(p, c->diagram, c->einst, c->coll, cl.temp_n, cl.temp_el, cl.el_conc, cl.h_conc,
                                               cl.ph2_conc, cl.oh2_conc, cl.he_conc);
                else if (start_mode == RT_START_LTE) boltzmann_populations(cl.temp_n, p, c->diagram);
                else memcpy(p, given.data(), sizeof(double) * N);
            };
            start();
            cs.calls = 0;
            bool found = ctl.calculate_populations(p, max_iter, min_err, accel != 0, 0);
            if (!found && accel && retry && name != "CH3OHa" && name != "CH3OHe") {
                start();
                cs.calls = 0;
                found = ctl.calculate_populations(p, max_iter_ext, min_err, false, 0);
            }
            if (st) {
                st[l].iters = cs.calls;
                st[l].converged = found ? 1 : 0;
                st[l].eq_error = ctl.eq_error;
                st[l].pop_error = ctl.pop_error;
                st[l].rel_error = ctl.rel_error;
            }
        }
        delete s;
    }
    if (seconds) *seconds = omp_get_wtime() - t0;
    delete c;
    return 0;
}

// The reference's calc_molecular_populations itself (warm-start chain + retry policy), verbosity=1 output parsed
// for the per-layer iteration counts (iteration table printed at iteration_control.h:217-229).
int ref_calc_molecular_populations(const rt_model_desc* d, const rt_layer_soa* lay, double* pop, int acceleration,
                                   rt_layer_status* st, double* seconds)
{
    ref_ctx* c = build_ctx(d, lay);
    std::stringbuf cap;
    const double t0 = omp_get_wtime();
    {
        cout_redirect r(&cap);
        calc_molecular_populations(c->cloud, c->scheme, c->diagram, c->einst, c->coll, pop, c->N, acceleration != 0, 1);
        std::cout.flush();
    }
    if (seconds) *seconds = omp_get_wtime() - t0;
    if (st) {
        for (int l = 0; l < c->L; l++) { st[l].iters = 0; st[l].converged = 1; st[l].eq_error = st[l].pop_error = st[l].rel_error = 0.; }
        std::istringstream in(cap.str());
        std::string line;
        int layer = -1;
        while (std::getline(in, line)) {
            size_t h = line.find("Level populations calculation:");
            if (h != std::string::npos) {
                // "<lay_nb> Level populations..." for a first attempt, bare header for the retry of the same layer
                std::istringstream pre(line.substr(0, h));
                int v;
                if (pre >> v) layer = v;
                if (layer >= 0 && layer < c->L) st[layer].iters = 0;
                continue;
            }
            if (line.compare(0, 5, "Can n") == 0) {
                // "Can not find solution for layers:" followed by the list
                std::string rest;
                if (std::getline(in, rest)) {
                    std::istringstream bl(rest);
                    int b;
                    while (bl >> b) if (b >= 0 && b < c->L) st[b].converged = 0;
                }
                break;
            }
            if (layer >= 0 && layer < c->L && !line.empty() && isdigit((unsigned char)line[0])) {
                std::istringstream ls(line);
                int it; double e1, e2, e3;
                if (ls >> it >> e1 >> e2 >> e3) { st[layer].iters = it; st[layer].eq_error = e1; st[layer].pop_error = e2; st[layer].rel_error = e3; }
            }
        }
    }
    delete c;
    return 0;
}

// transition_data_container::calc_inv / calc_gain / calc_line_profile for listed transitions.
// scalars [nb_trans][5] = {inv, gain, tau_eff, tau_max, lay_nb_hg}; nb_freq=300, nb_aspect=37, da=0.25,
// velocity_shift=5e5 are fixed inside the reference (transition_data.cpp:16-18,191).
int ref_maser_postpass(const rt_model_desc* d, const rt_layer_soa* lay, const double* pop, int nb_trans, const int* up,
                       const int* low, double* inv_arr, double* gain_arr, double* scalars, double* tau_vs_aspect,
                       double* tau_vs_freq)
{
    ref_ctx* c = build_ctx(d, lay);
    transition_data_container tc(c->cloud, c->diagram, c->einst);
    std::vector<double> p(pop, pop + (size_t)c->L * c->N);
    for (int t = 0; t < nb_trans; t++) {
        transition_data td(c->L, c->diagram->lev_array[low[t]], c->diagram->lev_array[up[t]]);
        tc.calc_inv(td, p.data());
        tc.calc_gain(td);
        tc.calc_line_profile(td);
        if (inv_arr) memcpy(inv_arr + (size_t)t * c->L, td.inv_arr, sizeof(double) * c->L);
        if (gain_arr) memcpy(gain_arr + (size_t)t * c->L, td.gain_arr, sizeof(double) * c->L);
        if (scalars) {
            double* s = scalars + 5 * t;
            s[0] = td.inv; s[1] = td.gain; s[2] = td.tau_eff; s[3] = td.tau_max; s[4] = td.lay_nb_hg;
        }
        if (tau_vs_aspect) memcpy(tau_vs_aspect + (size_t)t * td.nb_aspect_ratio, td.tau_vs_aspect_ratio, sizeof(double) * td.nb_aspect_ratio);
        if (tau_vs_freq) memcpy(tau_vs_freq + (size_t)t * td.nb_freq, td.tau_vs_frequency, sizeof(double) * td.nb_freq);
    }
    delete c;
    return 0;
}

// lim_luminosity_lvg (maser_luminosity.cpp:7-106) + calc_saturation_depth (transition_data.cpp:456-483) of the
// reference for listed transitions: the container is filled by transition_data_container::add (which also runs
// calc_inv / calc_gain / calc_line_profile / calc_exc_temp), entries are matched back by their level numbers.
// Per-layer outputs [nb_trans][L]; lum, tau_sat, temp_sat, lay_nb_hg [nb_trans].
int ref_lim_luminosity(const rt_model_desc* d, const rt_layer_soa* lay, const double* pop, int nb_trans, const int* up, const int* low,
                       double beaming_factor, double* lum_arr, double* pump_eff_arr, double* loss_rate_arr, double* pump_rate_arr,
                       double* emiss_coeff_arr, double* exc_temp_arr, double* lum, int* lay_nb_hg, double* tau_sat, double* temp_sat)
{
    ref_ctx* c = build_ctx(d, lay);
    null_buf nb;
    cout_redirect quiet(&nb);
    const int L = c->L;
    std::vector<double> p(pop, pop + (size_t)L * c->N);
    transition_data_container tc(c->cloud, c->diagram, c->einst);
    // what transition_data_container::add does per transition (transition_data.cpp:441-451), without its de-duplication by
    // quantum numbers (synthetic levels may share v, j, g) and without the final sort, so entry t stays transition t
    for (int t = 0; t < nb_trans; t++) {
        transition_data td(L, c->diagram->lev_array[low[t]], c->diagram->lev_array[up[t]]);
        tc.calc_inv(td, p.data());
        tc.calc_gain(td);
        tc.calc_line_profile(td);
        tc.calc_exc_temp(td, p.data());
        tc.data.push_back(td);
    }
    lim_luminosity_lvg(c->scheme, &tc, c->cloud, c->diagram, c->einst, c->coll, p.data());
    tc.calc_saturation_depth(beaming_factor);
    int found = 0;
    std::list<transition_data>::iterator it = tc.data.begin();
    for (int t = 0; t < nb_trans && it != tc.data.end(); t++, ++it) {
        if (it->trans->up_lev.nb != up[t] || it->trans->low_lev.nb != low[t]) break;
        found++;
        const size_t o = (size_t)t * L;
        if (lum_arr) memcpy(lum_arr + o, it->lum_arr, sizeof(double) * L);
        if (pump_eff_arr) memcpy(pump_eff_arr + o, it->pump_eff_arr, sizeof(double) * L);
        if (loss_rate_arr) memcpy(loss_rate_arr + o, it->loss_rate_arr, sizeof(double) * L);
        if (pump_rate_arr) memcpy(pump_rate_arr + o, it->pump_rate_arr, sizeof(double) * L);
        if (emiss_coeff_arr) memcpy(emiss_coeff_arr + o, it->emiss_coeff_arr, sizeof(double) * L);
        if (exc_temp_arr) memcpy(exc_temp_arr + o, it->exc_temp_arr, sizeof(double) * L);
        if (lum) lum[t] = it->lum;
        if (lay_nb_hg) lay_nb_hg[t] = it->lay_nb_hg;
        if (tau_sat) tau_sat[t] = it->tau_sat;
        if (temp_sat) temp_sat[t] = it->temp_sat;
    }
    delete c;
    return found == nb_trans ? 0 : 1;
}

} // extern "C"

// =========================================================================================================
// Adapter checks (include/rt_b200_adapter.hpp): built into oracle/_ref/libref_adapter.so, which links librt_b200.so.
// The reference objects are built from the flat problem as above; the ADAPTER then reads them back (through the
// reference's own class surface) and drives the CUDA library -- the drop-in path a maintainer would use.
// =========================================================================================================
#ifdef RT_WITH_ADAPTER
#include "../include/rt_b200_adapter.hpp"

extern "C" {

// flatten(reference objects) must reproduce the flat problem the objects were built from: returns the number of
// mismatching scalars/array elements (0 = byte-identical round trip), -1 on a structural mismatch.
int ref_adapter_roundtrip(const rt_model_desc* d, const rt_layer_soa* lay)
{
    ref_ctx* c = build_ctx(d, lay);
    rt_b200::flat_model fm;
    rt_b200::flat_cloud fc;
    rt_b200::flatten_model(c->scheme, c->diagram, c->einst, c->coll, d->species_rule, fm);
    rt_b200::flatten_cloud(c->cloud, fc);
    const rt_model_desc& q = fm.desc;
    long bad = 0;
    auto cmp = [&](const double* a, const double* b, size_t n) { for (size_t i = 0; i < n; i++) bad += (a[i] != b[i]); };
    if (q.nb_lev != d->nb_lev || q.nb_neutral_tables != d->nb_neutral_tables || q.nb_electron_tables != d->nb_electron_tables
        || q.scheme != d->scheme || q.dust_ncomp != d->dust_ncomp || q.lvg.nb_d != d->lvg.nb_d || q.lvg.nb_g != d->lvg.nb_g
        || fc.soa.nb_lay != lay->nb_lay || fc.soa.dust_ncomp != lay->dust_ncomp) { delete c; return -1; }
    const int N = d->nb_lev;
    cmp(q.energy, d->energy, N); cmp(q.einst, d->einst, (size_t)N * N);
    for (int i = 0; i < N; i++) bad += (q.g[i] != d->g[i]) + (d->v && q.v[i] != d->v[i]) + (d->j && q.j[i] != d->j[i]);
    bad += (q.mol_mass != d->mol_mass);
    for (int k = 0; k < q.nb_neutral_tables + q.nb_electron_tables; k++) {
        if (q.tables[k].nb_lev != d->tables[k].nb_lev || q.tables[k].jmax != d->tables[k].jmax || q.tables[k].interp != d->tables[k].interp) { delete c; return -1; }
        cmp(q.tables[k].tgrid, d->tables[k].tgrid, q.tables[k].jmax);
        cmp(q.tables[k].coeff, d->tables[k].coeff, (size_t)q.tables[k].jmax * q.tables[k].nb_lev * (q.tables[k].nb_lev - 1) / 2);
    }
    cmp(q.lvg.delta, d->lvg.delta, q.lvg.nb_d); cmp(q.lvg.gamma, d->lvg.gamma, q.lvg.nb_g);
    cmp(q.lvg.p, d->lvg.p, (size_t)q.lvg.nb_d * q.lvg.nb_g);
    if (d->scheme == RT_SCHEME_LINE_OVERLAP) {
        const rt_overlap_table* a[2] = {q.overlap1, q.overlap2};
        const rt_overlap_table* b[2] = {d->overlap1, d->overlap2};
        for (int t = 0; t < 2; t++) {
            if (a[t]->nb_d != b[t]->nb_d || a[t]->nb_dx != b[t]->nb_dx || a[t]->nb_gr != b[t]->nb_gr || a[t]->nb_g != b[t]->nb_g) { delete c; return -1; }
            for (int i = 0; i < a[t]->nb_d; i++) bad += (a[t]->delta[i] != log10(b[t]->delta[i]));
            cmp(a[t]->dx, b[t]->dx, a[t]->nb_dx); cmp(a[t]->gratio, b[t]->gratio, a[t]->nb_gr); cmp(a[t]->gamma, b[t]->gamma, a[t]->nb_g);
            cmp(a[t]->p, b[t]->p, (size_t)a[t]->nb_d * a[t]->nb_dx * a[t]->nb_gr * a[t]->nb_g);
        }
    }
    for (int k = 0; k < q.dust_ncomp; k++) {
        if (q.dust[k].nb_ph_en != d->dust[k].nb_ph_en) { delete c; return -1; }
        if (c->dust_exact) cmp(q.dust[k].ph_en, d->dust[k].ph_en, q.dust[k].nb_ph_en);
        cmp(q.dust[k].abs_coeff, d->dust[k].abs_coeff, q.dust[k].nb_ph_en);
        bad += (q.dust[k].wvl_exp != d->dust[k].wvl_exp);
    }
    const double* la[] = {fc.soa.temp_n, fc.soa.temp_el, fc.soa.el_conc, fc.soa.h_conc, fc.soa.ph2_conc, fc.soa.oh2_conc, fc.soa.he_conc,
                          fc.soa.mol_conc, fc.soa.vel_turb, fc.soa.velg_n, fc.soa.dz, fc.soa.vel_n};
    const double* lb[] = {lay->temp_n, lay->temp_el, lay->el_conc, lay->h_conc, lay->ph2_conc, lay->oh2_conc, lay->he_conc,
                          lay->mol_conc, lay->vel_turb, lay->velg_n, lay->dz, lay->vel_n};
    for (int k = 0; k < 12; k++) cmp(la[k], lb[k], lay->nb_lay);
    cmp(fc.soa.dust_conc, lay->dust_conc, (size_t)lay->nb_lay * lay->dust_ncomp);
    bad += (fc.soa.height != lay->height);
    delete c;
    return (int)bad;
}

// reference objects -> adapter -> librt_b200.so (GPU).  Returns 0 on success, 1 if the adapter reported failure.
int ref_adapter_calc_molecular_populations(const rt_model_desc* d, const rt_layer_soa* lay, double* pop, int acceleration,
                                           int start_mode, int device, rt_layer_status* st)
{
    ref_ctx* c = build_ctx(d, lay);
    null_buf nb;
    cout_redirect quiet(&nb);
    const bool ok = rt_b200::calc_molecular_populations_gpu(c->cloud, c->scheme, c->diagram, c->einst, c->coll, pop, c->N,
                                                            acceleration != 0, 0, start_mode, device, d->species_rule, st);
    delete c;
    return ok ? 0 : 1;
}


} // extern "C"

// ---- the per-call drop-in (rt_b200::scheme_gpu): the REFERENCE's controller drives GPU steps ---------------------------
namespace {
template <class S> struct counting_any {
    S* s;
    int calls;
    int get_vector_dim() { return s->get_vector_dim(); }
    void calc_new_pop(double* o, double* n, double& e) { calls++; s->calc_new_pop(o, n, e); }
};

template <class S> int gpu_scheme_solve(ref_ctx* c, const rt_model_desc* d, double* pop, int start_mode, int max_iter, double min_err,
                                        int accel, rt_layer_status* st, int lay_begin, int lay_end, int device)
{
    const int N = c->N;
    S s(c->tmp, 0, device, d->species_rule);
    s.set_dust_model(c->dust);
    s.init_molecule_data(c->diagram, c->einst, c->coll);
    counting_any<S> cs{&s, 0};
    iteration_control<counting_any<S> > ctl(&cs);
    for (int l = lay_begin; l < lay_end; l++) {
        const cloud_layer cl = c->cloud->lay_array[l];
        iteration_scheme_lvg* base = &s; // the setters are called the way radiative_transfer.cpp:238-242 calls them
        base->set_vel_grad(cl.velg_n);
        std::vector<double> dc(cl.dust_grain_conc), dt(cl.dust_grain_temp);
        base->set_dust_parameters(dc, dt);
        base->set_parameters(cl.temp_n, cl.temp_el, cl.el_conc, cl.h_conc, cl.ph2_conc, cl.oh2_conc, cl.he_conc, cl.mol_conc, cl.vel_turb);
        double* p = pop + (size_t)l * N;
        if (start_mode == RT_START_BOUNDARY)
            boundary_layer_populations(p, c->diagram, c->einst, c->coll, cl.temp_n, cl.temp_el, cl.el_conc, cl.h_conc, cl.ph2_conc,
                                       cl.oh2_conc, cl.he_conc);
        else if (start_mode == RT_START_LTE) boltzmann_populations(cl.temp_n, p, c->diagram);
        cs.calls = 0;
        const bool found = ctl.calculate_populations(p, max_iter, min_err, accel != 0, 0);
        if (st) {
            st[l].iters = cs.calls; st[l].converged = found ? 1 : 0;
            st[l].eq_error = ctl.eq_error; st[l].pop_error = ctl.pop_error; st[l].rel_error = ctl.rel_error;
        }
    }
    return 0;
}

// operator() and intensity_calc of the GPU scheme called through a BASE pointer (virtual dispatch), layer l
template <class S> int gpu_scheme_probe(ref_ctx* c, const rt_model_desc* d, int l, const double* pop, double* matrix, double* df,
                                        double* new_pop, double* eq_err, int n, const int* up, const int* low, double* J, int device)
{
    const int N = c->N;
    S s(c->tmp, 0, device, d->species_rule);
    iteration_scheme_lvg* base = &s;
    base->set_dust_model(c->dust);
    base->init_molecule_data(c->diagram, c->einst, c->coll);
    const cloud_layer cl = c->cloud->lay_array[l];
    base->set_vel_grad(cl.velg_n);
    std::vector<double> dc(cl.dust_grain_conc), dt(cl.dust_grain_temp);
    base->set_dust_parameters(dc, dt);
    base->set_parameters(cl.temp_n, cl.temp_el, cl.el_conc, cl.h_conc, cl.ph2_conc, cl.oh2_conc, cl.he_conc, cl.mol_conc, cl.vel_turb);
    std::vector<double> p(pop, pop + N);
    (*base)(N, p.data(), df);
    struct peek : public S { using S::S; double** mat() { return this->matrix; } };
    memcpy(matrix, *static_cast<peek*>(static_cast<S*>(base))->mat(), sizeof(double) * N * N);
    // base-class calc_new_pop: GPU assembly through the virtual operator(), then the reference's own LU
    base->calc_new_pop(p.data(), new_pop, *eq_err);
    for (int i = 0; i < n; i++) base->intensity_calc(up[i], low[i], p.data(), J[i]);
    return 0;
}
} // namespace

extern "C" {

int ref_adapter_scheme_solve(const rt_model_desc* d, const rt_layer_soa* lay, double* pop, int start_mode, int max_iter, double min_err,
                             int accel, rt_layer_status* st, int lay_begin, int lay_end, int device)
{
    ref_ctx* c = build_ctx(d, lay);
    if (lay_end < 0 || lay_end > c->L) lay_end = c->L;
    null_buf nb;
    cout_redirect quiet(&nb);
    const int rc = c->overlap
        ? gpu_scheme_solve<rt_b200::iteration_scheme_line_overlap_gpu>(c, d, pop, start_mode, max_iter, min_err, accel, st, lay_begin, lay_end, device)
        : gpu_scheme_solve<rt_b200::iteration_scheme_lvg_gpu>(c, d, pop, start_mode, max_iter, min_err, accel, st, lay_begin, lay_end, device);
    delete c;
    return rc;
}

// lim_luminosity_lvg_gpu on the reference's container, then the reference's own calc_saturation_depth on the result
int ref_adapter_lim_luminosity(const rt_model_desc* d, const rt_layer_soa* lay, const double* pop, int nb_trans, const int* up,
                               const int* low, double beaming_factor, double* lum_arr, double* loss_rate_arr, double* pump_rate_arr,
                               double* lum, double* tau_sat, int device)
{
    ref_ctx* c = build_ctx(d, lay);
    null_buf nb;
    cout_redirect quiet(&nb);
    const int L = c->L;
    std::vector<double> p(pop, pop + (size_t)L * c->N);
    transition_data_container tc(c->cloud, c->diagram, c->einst);
    for (int t = 0; t < nb_trans; t++) {
        transition_data td(L, c->diagram->lev_array[low[t]], c->diagram->lev_array[up[t]]);
        tc.calc_inv(td, p.data());
        tc.calc_gain(td);
        tc.calc_exc_temp(td, p.data());
        tc.data.push_back(td);
    }
    const bool ok = rt_b200::lim_luminosity_lvg_gpu(c->scheme, &tc, c->cloud, c->diagram, c->einst, c->coll, p.data(), device, d->species_rule);
    tc.calc_saturation_depth(beaming_factor);
    int t = 0;
    for (std::list<transition_data>::iterator it = tc.data.begin(); it != tc.data.end(); ++it, ++t) {
        memcpy(lum_arr + (size_t)t * L, it->lum_arr, sizeof(double) * L);
        memcpy(loss_rate_arr + (size_t)t * L, it->loss_rate_arr, sizeof(double) * L);
        memcpy(pump_rate_arr + (size_t)t * L, it->pump_rate_arr, sizeof(double) * L);
        lum[t] = it->lum; tau_sat[t] = it->tau_sat;
    }
    delete c;
    return ok ? 0 : 1;
}

int ref_adapter_scheme_probe(const rt_model_desc* d, const rt_layer_soa* lay, int l, const double* pop, double* matrix, double* df,
                             double* new_pop, double* eq_err, int n, const int* up, const int* low, double* J, int device)
{
    ref_ctx* c = build_ctx(d, lay);
    null_buf nb;
    cout_redirect quiet(&nb);
    const int rc = c->overlap
        ? gpu_scheme_probe<rt_b200::iteration_scheme_line_overlap_gpu>(c, d, l, pop, matrix, df, new_pop, eq_err, n, up, low, J, device)
        : gpu_scheme_probe<rt_b200::iteration_scheme_lvg_gpu>(c, d, l, pop, matrix, df, new_pop, eq_err, n, up, low, J, device);
    delete c;
    return rc;
}

} // extern "C"
#endif
