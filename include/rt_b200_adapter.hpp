// rt_b200_adapter.hpp -- header-only C++ adapter between the REFERENCE's own classes and the C ABI of librt_b200.so.
//
// A maintainer of nesterenok/radiative_transfer adds this header to the driver and replaces the call
//     calc_molecular_populations(cloud, it_scheme, levels, einst, coll, popul, nb_lev, accel, verbosity)
// (radiative_transfer.cpp:219, call sites :440,571,670,797,925,1058,...) by
//     rt_b200::calc_molecular_populations_gpu(cloud, it_scheme, levels, einst, coll, popul, nb_lev, accel, verbosity)
// -- same arguments, same in/out array `popul[lay*nb_lev + lev]`, same stdout report of bad layers.
// Everything is read from the reference's objects (no text files are re-parsed): protected members are reached
// through derived accessor classes (pointer-to-member of a protected base member, legal C++).
// Compiles only inside the reference's tree (it includes the reference's headers); nothing here is used by the
// Python host side.  Exercised by oracle/ref_harness.cpp (ref_adapter_*), tests/test_adapter.py.
#pragma once
#include <cmath>
#include <cstring>
#include <iostream>
#include <list>
#include <string>
#include <vector>

#include "cloud_data.h"
#include "coll_rates.h"
#include "coll_rates_ch3oh.h"
#include "coll_rates_h2co.h"
#include "coll_rates_h2o.h"
#include "coll_rates_nh3.h"
#include "coll_rates_oh.h"
#include "dust_model.h"
#include "iteration_lvg.h"
#include "lvg_method_functions.h"
#include "spectroscopy.h"
#include "transition_data.h"

#include "rt_b200.h"

namespace rt_b200 {

// ---- accessors to protected members of the reference's classes ------------------------------------------------
struct coll_access : public collisional_transitions {
    static const std::vector<collision_data*>& data(const collisional_transitions* c) { return c->*(&coll_access::coll_data); }
    static int n1(const collisional_transitions* c) { return c->*(&coll_access::nb1); }
    static int n2(const collisional_transitions* c) { return c->*(&coll_access::nb2); }
};
struct lvg_access : public lvg_method_data {
    static int nd(const lvg_method_data* t) { return t->*(&lvg_access::nb_d); }
    static int ng(const lvg_method_data* t) { return t->*(&lvg_access::nb_g); }
    static const double* delta(const lvg_method_data* t) { return t->*(&lvg_access::delta_arr); }
    static const double* gamma(const lvg_method_data* t) { return t->*(&lvg_access::gamma_arr); }
    static double** p(const lvg_method_data* t) { return t->*(&lvg_access::p_arr); }
};
struct ovl_access : public lvg_line_overlap_data {
    static int nd(const lvg_line_overlap_data* t) { return t->*(&ovl_access::nb_d); }
    static int ng(const lvg_line_overlap_data* t) { return t->*(&ovl_access::nb_g); }
    static int ngr(const lvg_line_overlap_data* t) { return t->*(&ovl_access::nb_gr); }
    static int ndx(const lvg_line_overlap_data* t) { return t->*(&ovl_access::nb_dx); }
    static const double* delta(const lvg_line_overlap_data* t) { return t->*(&ovl_access::delta_arr); } // log10 (cpp:293)
    static const double* gamma(const lvg_line_overlap_data* t) { return t->*(&ovl_access::gamma_arr); }
    static const double* gratio(const lvg_line_overlap_data* t) { return t->*(&ovl_access::gratio_arr); }
    static const double* dx(const lvg_line_overlap_data* t) { return t->*(&ovl_access::dx_arr); }
    static double** p(const lvg_line_overlap_data* t) { return t->*(&ovl_access::p_arr); }
};
struct dustc_access : public dust_component {
    static int n(const dust_component* c) { return c->*(&dustc_access::nb_ph_en); }
    static const double* en(const dust_component* c) { return c->*(&dustc_access::ph_en_arr); }
    static const double* abs(const dust_component* c) { return c->*(&dustc_access::abs_coeff); }
};
struct scheme_access : public iteration_scheme_lvg {
    static const dust_model* dust_of(const iteration_scheme_lvg* s) { return s->*(&scheme_access::dust); }
    static const lvg_method_data* lvg_of(const iteration_scheme_lvg* s) { return s->*(&scheme_access::loss_func_line_phot); }
};
struct overlap_access : public iteration_scheme_line_overlap {
    static const std::vector<hfs_lines>& groups(const iteration_scheme_line_overlap* s) { return s->*(&overlap_access::line_list); }
    static const lvg_line_overlap_data* t1(const iteration_scheme_line_overlap* s) { return s->*(&overlap_access::line_overlap1); }
    static const lvg_line_overlap_data* t2(const iteration_scheme_line_overlap* s) { return s->*(&overlap_access::line_overlap2); }
    static double maxdx(const iteration_scheme_line_overlap* s) { return s->*(&overlap_access::max_dx); }
};

// the species mixing rule implemented by the dynamic type of `coll` (coll_rates_ch3oh.cpp:484, coll_rates_oh.cpp:392,
// coll_rates_h2o.cpp:530); anything else follows the base-class rule coll_rates.cpp:181-197
inline int species_rule_of(const collisional_transitions* coll)
{
    if (dynamic_cast<const ch3oh_collisions*>(coll)) return RT_RULE_CH3OH;
    if (dynamic_cast<const oh_hf_collisions*>(coll)) return RT_RULE_OH_HF;
    if (dynamic_cast<const h2o_collisions*>(coll)) return RT_RULE_H2O;
    if (dynamic_cast<const nh3_collisions*>(coll)) return RT_RULE_NH3;
    if (dynamic_cast<const h2co_collisions*>(coll)) return RT_RULE_H2CO;
    return RT_RULE_GENERIC;
}

// ---- flattened copies (own the storage the rt_*_desc structs point into) ------------------------------------------
struct flat_model {
    std::vector<double> energy, j, einst, lvg_delta, lvg_gamma, lvg_p;
    std::vector<int> g, v, groups;
    std::vector<std::vector<double>> tgrid, coeff, dust_en, dust_abs;
    std::vector<rt_coll_table> tables;
    std::vector<rt_dust_comp> dust;
    std::vector<double> ov_p[2];
    rt_overlap_table ov[2];
    std::string name;
    rt_model_desc desc;
};

inline void flatten_overlap(const lvg_line_overlap_data* t, std::vector<double>& p, rt_overlap_table& o)
{
    o.nb_d = ovl_access::nd(t); o.nb_dx = ovl_access::ndx(t); o.nb_gr = ovl_access::ngr(t); o.nb_g = ovl_access::ng(t);
    o.delta = ovl_access::delta(t); o.delta_is_log10 = 1; // the reference stores log10(delta), lvg_method_functions.cpp:293
    o.dx = ovl_access::dx(t); o.gratio = ovl_access::gratio(t); o.gamma = ovl_access::gamma(t);
    const size_t rows = (size_t)o.nb_d * o.nb_dx, cols = (size_t)o.nb_gr * o.nb_g;
    p.resize(rows * cols);
    for (size_t r = 0; r < rows; r++) memcpy(&p[r * cols], ovl_access::p(t)[r], sizeof(double) * cols);
    o.p = p.data();
    o.reserved0 = 0;
}

// Reads energy_diagram / einstein_coeff / collisional_transitions and, through the scheme, the dust model and the
// escape-probability tables (for iteration_scheme_line_overlap also the reference's own line_list and max_dx).
inline void flatten_model(const iteration_scheme_lvg* scheme, const energy_diagram* levels, const einstein_coeff* einst,
                          const collisional_transitions* coll, int species_rule, flat_model& f)
{
    const int N = levels->nb_lev;
    rt_model_desc& d = f.desc;
    memset(&d, 0, sizeof d);
    rt_params_default(&d.params);
    f.energy.resize(N); f.j.resize(N); f.g.resize(N); f.v.resize(N);
    for (int i = 0; i < N; i++) {
        const energy_level& l = levels->lev_array[i];
        f.energy[i] = l.energy; f.g[i] = l.g; f.v[i] = l.v; f.j[i] = l.j;
    }
    f.einst.resize((size_t)N * N);
    for (int i = 0; i < N; i++) memcpy(&f.einst[(size_t)i * N], einst->arr[i], sizeof(double) * N);
    f.name = levels->mol.name;
    d.nb_lev = N; d.species_rule = species_rule < 0 ? species_rule_of(coll) : species_rule;
    d.energy = f.energy.data(); d.g = f.g.data(); d.v = f.v.data(); d.j = f.j.data();
    d.mol_mass = levels->mol.mass; d.mol_name = f.name.c_str(); d.einst = f.einst.data();

    const std::vector<collision_data*>& cd = coll_access::data(coll);
    const int nb1 = coll_access::n1(coll), nb2 = coll_access::n2(coll); // tables [nb2, nb3) are ion data: not used by the scheme
    f.tables.resize(nb2); f.tgrid.resize(nb2); f.coeff.resize(nb2);
    for (int k = 0; k < nb2; k++) {
        const collision_data* t = cd[k];
        f.tgrid[k].assign(t->tgrid, t->tgrid + t->jmax);
        f.coeff[k].resize((size_t)t->imax * t->jmax);
        for (int i = 0; i < t->imax; i++) memcpy(&f.coeff[k][(size_t)i * t->jmax], t->coeff[i], sizeof(double) * t->jmax);
        f.tables[k].nb_lev = t->nb_lev; f.tables[k].jmax = t->jmax;
        f.tables[k].tgrid = f.tgrid[k].data(); f.tables[k].coeff = f.coeff[k].data();
        // a table object of the cubic-spline class (coll_rates.h:44-53) is evaluated with its own rule on the GPU
        f.tables[k].interp = dynamic_cast<const collision_data_cub_spline*>(t) ? RT_INTERP_CUBIC_SPLINE : RT_INTERP_LINEAR;
        f.tables[k].reserved0 = 0;
    }
    d.nb_neutral_tables = nb1; d.nb_electron_tables = nb2 - nb1; d.tables = f.tables.data();

    const lvg_method_data* lt = scheme_access::lvg_of(scheme);
    const int nd = lvg_access::nd(lt), ng = lvg_access::ng(lt);
    f.lvg_delta.assign(lvg_access::delta(lt), lvg_access::delta(lt) + nd);
    f.lvg_gamma.assign(lvg_access::gamma(lt), lvg_access::gamma(lt) + ng);
    f.lvg_p.resize((size_t)nd * ng);
    for (int i = 0; i < nd; i++) memcpy(&f.lvg_p[(size_t)i * ng], lvg_access::p(lt)[i], sizeof(double) * ng);
    d.lvg.nb_d = nd; d.lvg.nb_g = ng; d.lvg.delta = f.lvg_delta.data(); d.lvg.gamma = f.lvg_gamma.data(); d.lvg.p = f.lvg_p.data();

    d.scheme = RT_SCHEME_LVG;
    if (const iteration_scheme_line_overlap* so = dynamic_cast<const iteration_scheme_line_overlap*>(scheme)) {
        d.scheme = RT_SCHEME_LINE_OVERLAP;
        d.params.max_dx = overlap_access::maxdx(so);
        for (const hfs_lines& h : overlap_access::groups(so)) { // the reference's own grouping, iteration_lvg.cpp:317-346
            const int g5[5] = {h.nb, h.upl[0], h.lowl[0], h.nb > 1 ? h.upl[1] : -1, h.nb > 1 ? h.lowl[1] : -1};
            f.groups.insert(f.groups.end(), g5, g5 + 5);
        }
        d.nb_line_groups = (int)f.groups.size() / 5; d.line_groups = f.groups.data();
        flatten_overlap(overlap_access::t1(so), f.ov_p[0], f.ov[0]);
        flatten_overlap(overlap_access::t2(so), f.ov_p[1], f.ov[1]);
        d.overlap1 = &f.ov[0]; d.overlap2 = &f.ov[1];
    }

    const dust_model* dm = scheme_access::dust_of(scheme);
    f.dust.resize(dm->nb_of_comp); f.dust_en.resize(dm->nb_of_comp); f.dust_abs.resize(dm->nb_of_comp);
    for (int c = 0; c < dm->nb_of_comp; c++) {
        const dust_component* dc = dm->components[c];
        const int n = dustc_access::n(dc);
        f.dust_en[c].assign(dustc_access::en(dc), dustc_access::en(dc) + n);
        f.dust_abs[c].assign(dustc_access::abs(dc), dustc_access::abs(dc) + n);
        f.dust[c].nb_ph_en = n; f.dust[c].reserved0 = 0; f.dust[c].ph_en = f.dust_en[c].data();
        f.dust[c].abs_coeff = f.dust_abs[c].data(); f.dust[c].wvl_exp = dc->wvl_exp;
    }
    d.dust_ncomp = dm->nb_of_comp; d.dust = f.dust.data();
}

// The oH2O 6_16 -> 5_23 (22.2 GHz) line whose gain transition_data_container::calc_gain evaluates with a widened
// thermal width (transition_data.cpp:249-260): found by the quantum numbers the reference tests for and handed to the
// library by level index (rt_set_gain_width_line), once per model.  Returns false (and clears the setting) for any
// other molecule.
inline bool set_h2o_22ghz_line(rt_model* m, const energy_diagram* levels)
{
    int up = -1, low = -1;
    if (levels->mol.name == "oH2O" && levels->mol.isotop == 1) {
        for (int i = 0; i < levels->nb_lev; i++) {
            const energy_level& lv = levels->lev_array[i];
            if (lv.v == 0 && lv.j == 6 && lv.k1 == 1 && lv.k2 == 6) up = i;
            if (lv.j == 5 && lv.k1 == 2 && lv.k2 == 3 && low < 0) low = i;
        }
    }
    const bool found = up >= 0 && low >= 0 && low < up;
    rt_set_gain_width_line(m, found ? up : -1, low, 5.e+4);
    return found;
}

struct flat_cloud {
    std::vector<double> a[12], dust_conc;
    rt_layer_soa soa;
};

// cloud_data / cloud_layer (cloud_data.h:24-78) -> layer-major SoA
inline void flatten_cloud(const cloud_data* cloud, flat_cloud& f)
{
    const int L = cloud->nb_lay;
    const int nc = L > 0 ? (int)cloud->lay_array[0].dust_grain_conc.size() : 0;
    for (auto& v : f.a) v.resize(L);
    f.dust_conc.resize((size_t)L * nc);
    for (int l = 0; l < L; l++) {
        const cloud_layer& c = cloud->lay_array[l];
        const double v[12] = {c.temp_n, c.temp_el, c.el_conc, c.h_conc, c.ph2_conc, c.oh2_conc, c.he_conc, c.mol_conc,
                              c.vel_turb, c.velg_n, c.dz, c.vel_n};
        for (int k = 0; k < 12; k++) f.a[k][l] = v[k];
        for (int k = 0; k < nc; k++) f.dust_conc[(size_t)l * nc + k] = k < (int)c.dust_grain_conc.size() ? c.dust_grain_conc[k] : 0.;
    }
    rt_layer_soa& s = f.soa;
    memset(&s, 0, sizeof s);
    s.nb_lay = L; s.dust_ncomp = nc;
    s.temp_n = f.a[0].data(); s.temp_el = f.a[1].data(); s.el_conc = f.a[2].data(); s.h_conc = f.a[3].data();
    s.ph2_conc = f.a[4].data(); s.oh2_conc = f.a[5].data(); s.he_conc = f.a[6].data(); s.mol_conc = f.a[7].data();
    s.vel_turb = f.a[8].data(); s.velg_n = f.a[9].data(); s.dz = f.a[10].data(); s.vel_n = f.a[11].data();
    s.dust_conc = f.dust_conc.data();
    s.height = cloud->get_height();
}

// Drop-in for calc_molecular_populations (radiative_transfer.cpp:219-289).  The DEFAULT start_mode RT_START_CHAINED is the
// reference's own semantics: layer l starts from the answer of layer l-1 when that one converged (:247-249), else from
// boundary_layer_populations (:250-253); iterates, iteration counts and convergence flags are those of the unmodified
// driver.  A single cloud is then a serial chain of one-layer solves (a sweep batches the chains of many clouds: stack
// them with rt_layer_soa::model_first).  start_mode = RT_START_BOUNDARY is the explicit opt-in to the batch-parallel
// independent start -- every layer from boundary_layer_populations, all layers at once on the GPU; same fixed points, but
// NOT the driver's iterates: a layer with several solutions, or one that does not converge, can end elsewhere.
// Fatal conditions are reported like the reference does (message to cout) but return false instead of exit(1).
// max_iter_acc / max_iter_ext / min_err: MAX_NB_ITER_ACC, MAX_NB_ITER_EXT (:26-27), rel_population_error (:45).
inline bool calc_molecular_populations_gpu(cloud_data* cloud, iteration_scheme_lvg* it_scheme_lvg, energy_diagram* mol_levels,
                                           einstein_coeff* mol_einst, collisional_transitions* mol_coll, double* mol_popul,
                                           int nb_lev, bool acceleration, int verbosity, int start_mode = RT_START_CHAINED,
                                           int device = 0, int species_rule = -1, rt_layer_status* status = nullptr,
                                           int max_iter_acc = 150, int max_iter_ext = 15000, double min_err = 1.e-5)
{
    flat_model fm;
    flat_cloud fc;
    flatten_model(it_scheme_lvg, mol_levels, mol_einst, mol_coll, species_rule, fm);
    flatten_cloud(cloud, fc);
    if (nb_lev != fm.desc.nb_lev) { std::cout << "rt_b200: nb_lev differs from the energy diagram" << std::endl; return false; }
    rt_model* m = nullptr;
    rt_status rc = rt_model_create(&fm.desc, device, &m);
    if (rc != RT_OK) { std::cout << "rt_b200: " << rt_status_string(rc) << ": " << rt_last_error(nullptr) << std::endl; return false; }
    std::vector<rt_layer_status> st(cloud->nb_lay);
    rc = rt_cloud_upload(m, &fc.soa);
    if (rc == RT_OK)
        rc = rt_calc_molecular_populations(m, mol_popul, start_mode, acceleration ? 1 : 0, max_iter_acc, max_iter_ext, min_err, st.data());
    if (rc != RT_OK) std::cout << "rt_b200: " << rt_status_string(rc) << ": " << rt_last_error(m) << std::endl;
    rt_model_destroy(m);
    if (rc != RT_OK) return false;
    bool any_bad = false;
    for (int l = 0; l < cloud->nb_lay; l++) any_bad = any_bad || !st[l].converged;
    if (verbosity && any_bad) { // radiative_transfer.cpp:283-288
        std::cout << "Can not find solution for layers:" << std::endl;
        for (int l = 0; l < cloud->nb_lay; l++)
            if (!st[l].converged) std::cout << l << " ";
        std::cout << std::endl;
    }
    if (status) memcpy(status, st.data(), sizeof(rt_layer_status) * st.size());
    return true;
}

// Drop-in for lim_luminosity_lvg (maser_luminosity.cpp:7-106): same arguments; fills lum, lum_arr, pump_eff_arr,
// loss_rate_arr, pump_rate_arr, emiss_coeff_arr of every entry of container->data.  literal_layer0 = true keeps the
// reference's behaviour of evaluating J of the loss terms with the populations of layer 0 (see rt_lim_luminosity).
// calc_saturation_depth of the container can be called afterwards exactly as before.
inline bool lim_luminosity_lvg_gpu(iteration_scheme_lvg* calc_scheme, transition_data_container* container, const cloud_data* cloud,
                                   const energy_diagram* diagram, const einstein_coeff* einst_coeff, collisional_transitions* coll_trans,
                                   double* level_pop, int device = 0, int species_rule = -1, bool literal_layer0 = true)
{
    const int L = cloud->nb_lay, T = (int)container->data.size();
    if (T == 0) return true;
    flat_model fm;
    flat_cloud fc;
    flatten_model(calc_scheme, diagram, einst_coeff, coll_trans, species_rule, fm);
    flatten_cloud(cloud, fc);
    rt_model* m = nullptr;
    rt_status rc = rt_model_create(&fm.desc, device, &m);
    if (rc != RT_OK) { std::cout << "rt_b200: " << rt_status_string(rc) << ": " << rt_last_error(nullptr) << std::endl; return false; }
    std::vector<int> up(T), low(T);
    int t = 0;
    for (std::list<transition_data>::iterator it = container->data.begin(); it != container->data.end(); ++it, ++t) {
        up[t] = it->trans->up_lev.nb; low[t] = it->trans->low_lev.nb;
    }
    std::vector<double> a[5], lum(T);
    for (auto& v : a) v.resize((size_t)T * L);
    rc = rt_cloud_upload(m, &fc.soa);
    if (rc == RT_OK)
        rc = rt_lim_luminosity(m, level_pop, T, up.data(), low.data(), literal_layer0 ? 1 : 0, a[0].data(), a[1].data(), a[2].data(),
                               a[3].data(), a[4].data(), nullptr, lum.data(), nullptr, 1., nullptr, nullptr);
    if (rc != RT_OK) std::cout << "rt_b200: " << rt_status_string(rc) << ": " << rt_last_error(m) << std::endl;
    rt_model_destroy(m);
    if (rc != RT_OK) return false;
    t = 0;
    for (std::list<transition_data>::iterator it = container->data.begin(); it != container->data.end(); ++it, ++t) {
        const size_t o = (size_t)t * L;
        memcpy(it->lum_arr, &a[0][o], sizeof(double) * L);
        memcpy(it->pump_eff_arr, &a[1][o], sizeof(double) * L);
        memcpy(it->loss_rate_arr, &a[2][o], sizeof(double) * L);
        memcpy(it->pump_rate_arr, &a[3][o], sizeof(double) * L);
        memcpy(it->emiss_coeff_arr, &a[4][o], sizeof(double) * L);
        it->lum = lum[t];
    }
    return true;
}

// ---------------------------------------------------------------------------------------------------------------
// Per-call drop-in: a scheme object whose step runs on the GPU.  It IS an iteration_scheme_lvg (or
// iteration_scheme_line_overlap), so the reference's three per-layer setters, its `iteration_control<T>` controller
// (duck-typed on get_vector_dim() / calc_new_pop(), iteration_control.h:101,210) and lim_luminosity_lvg
// (intensity_calc, maser_luminosity.cpp:54-58) keep working on it unchanged:
//
//     iteration_scheme_lvg_gpu scheme(data_path, verbosity);          // instead of iteration_scheme_lvg
//     scheme.set_dust_model(dust);  scheme.init_molecule_data(levels, einst, coll);
//     iteration_control<iteration_scheme_lvg_gpu> ctl(&scheme);       // the reference's controller, GPU steps
//     scheme.set_vel_grad(..); scheme.set_dust_parameters(..); scheme.set_parameters(..);   // radiative_transfer.cpp:238-242
//     ctl.calculate_populations(pop, max_iter, accel);
//
// The layer state is PULLED from the base object's protected members at every call (temp_n, temp_el, mol_conc, vel_width,
// vel_grad, coll_partn_conc, dgrain_conc), so it does not matter through which pointer type the setters were called;
// vel_width and the partner concentrations go through rt_layer_soa::vel_width / partner_conc verbatim.
// calc_new_pop is not virtual in the reference (iteration_lvg.h:53): the GPU step is taken when the controller is
// instantiated on this type; through a base pointer the virtual operator() below still assembles on the GPU and the
// base class's own LU finishes the step.  One layer per call: this is the fidelity path (step-by-step comparison with
// the CPU classes), the throughput path is calc_molecular_populations_gpu above.
template <class Base> class scheme_gpu : public Base
{
    rt_model* model_ = nullptr;
    int device_, rule_;
    flat_model fm_;
    std::vector<double> state_; // what the resident 1-layer cloud was built from

    bool sync()
    {
        if (!model_) {
            flatten_model(this, this->diagram, this->einst_coeff, this->coll_trans, rule_, fm_);
            rt_status rc = rt_model_create(&fm_.desc, device_, &model_);
            if (rc != RT_OK) { std::cout << "rt_b200: " << rt_status_string(rc) << ": " << rt_last_error(nullptr) << std::endl; model_ = nullptr; return false; }
            state_.clear();
        }
        const int K = fm_.desc.nb_neutral_tables + fm_.desc.nb_electron_tables;
        std::vector<double> st = {this->temp_n, this->temp_el, this->mol_conc, this->vel_width, this->vel_grad};
        for (int k = 0; k < K; k++) st.push_back(this->coll_partn_conc ? this->coll_partn_conc[k] : 0.);
        st.insert(st.end(), this->dgrain_conc.begin(), this->dgrain_conc.end());
        bool same = st.size() == state_.size();
        for (size_t i = 0; same && i < st.size(); i++) same = memcmp(&st[i], &state_[i], sizeof(double)) == 0;
        if (same) return true;
        const double zero = 0.;
        // vel_turb only feeds the post-pass; recovered from vel_width for completeness (iteration_lvg.cpp:65 inverted)
        const double vt2 = this->vel_width * this->vel_width - 2. * fm_.desc.params.boltzmann_constant * this->temp_n / fm_.desc.mol_mass;
        const double vt = vt2 > 0. ? sqrt(vt2) : 0.;
        rt_layer_soa s;
        memset(&s, 0, sizeof s);
        s.nb_lay = 1; s.dust_ncomp = (int)this->dgrain_conc.size();
        s.temp_n = &st[0]; s.temp_el = &st[1]; s.mol_conc = &st[2]; s.vel_width = &st[3]; s.velg_n = &st[4];
        s.el_conc = s.h_conc = s.ph2_conc = s.oh2_conc = s.he_conc = &zero; // superseded by partner_conc
        s.partner_conc = st.data() + 5;
        s.vel_turb = &vt;
        s.dust_conc = st.data() + 5 + K;
        rt_status rc = rt_cloud_upload(model_, &s);
        if (rc != RT_OK) { std::cout << "rt_b200: " << rt_status_string(rc) << ": " << rt_last_error(model_) << std::endl; state_.clear(); return false; }
        state_.swap(st);
        return true;
    }
    void fail(const char* what) const
    {
        std::cout << "rt_b200: " << what << " failed: " << rt_last_error(model_) << std::endl;
        exit(1); // the reference's own reaction to a fatal condition (iteration_lvg.cpp:82-83)
    }

public:
    scheme_gpu(const std::string& data_path, int verbosity = 1, int device = 0, int species_rule = -1)
        : Base(data_path, verbosity), device_(device), rule_(species_rule) {}
    ~scheme_gpu() { if (model_) rt_model_destroy(model_); }
    scheme_gpu(const scheme_gpu&) = delete;
    scheme_gpu& operator=(const scheme_gpu&) = delete;

    void init_molecule_data(const energy_diagram* d, const einstein_coeff* e, const collisional_transitions* c)
    {
        Base::init_molecule_data(d, e, c);
        if (model_) { rt_model_destroy(model_); model_ = nullptr; } // molecule changed: rebuilt at the next call
    }
    // == iteration_scheme_lvg::calc_new_pop (iteration_lvg.cpp:87-110): assembly + LU solve on the GPU
    void calc_new_pop(double* old_pop, double* new_pop, double& eq_error)
    {
        if (!sync() || rt_step(model_, old_pop, new_pop, &eq_error) != RT_OK) fail("rt_step");
    }
    // == operator() (iteration_lvg.cpp:112-161, :348-426): matrix into the base object's `matrix`, residual into df
    void operator()(int dim, double* level_pop, double* df)
    {
        if (dim != this->nb_mol_lev || !sync() || rt_assemble(model_, level_pop, *this->matrix, df) != RT_OK) fail("rt_assemble");
    }
    // == intensity_calc (iteration_lvg.cpp:163-185; overlap scheme :503-): dimensionless mean intensity of one line
    void intensity_calc(int upl, int lowl, double* lev_pop, double& intensity) const
    {
        scheme_gpu* self = const_cast<scheme_gpu*>(this);
        if (!self->sync() || rt_intensity(model_, 0, upl, lowl, lev_pop, &intensity) != RT_OK) fail("rt_intensity");
    }
    rt_model* model() { return sync() ? model_ : nullptr; }
};
typedef scheme_gpu<iteration_scheme_lvg> iteration_scheme_lvg_gpu;
typedef scheme_gpu<iteration_scheme_line_overlap> iteration_scheme_line_overlap_gpu;

} // namespace rt_b200
