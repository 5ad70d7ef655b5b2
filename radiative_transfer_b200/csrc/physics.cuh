// physics.cuh -- device functions shared by the coupling kernels: table searches, escape-probability
// interpolation, per-line mean intensity.  Each cites the reference lines whose arithmetic it follows
// (paths relative to /root/reference/radiative_transfer/).
#pragma once
#include "rt_internal.h"

// locate_index of the reference's missing interpolation library (contract: oracle/shim/utils.h):
// arr[k] <= x < arr[k+1]; -1 below the grid, n-1 at or above the last node.
__device__ __forceinline__ int dev_locate_index(const double* __restrict__ arr, int n, double x)
{
    if (!(x >= arr[0])) return -1;
    if (x >= arr[n - 1]) return n - 1;
    int lo = 0, hi = n - 1;
    while (hi - lo > 1) {
        const int mid = lo + ((hi - lo) >> 1);
        if (arr[mid] <= x) lo = mid; else hi = mid;
    }
    return lo;
}

// lvg_method_data::get_esc_func, lvg_method_functions.cpp:74-110: bilinear in linear (delta, gamma), edge clamp
// with t/u in {0,1}, result clamped to [0,1].
__device__ __forceinline__ double dev_esc_func(const DevLvg& t, double gamma, double delta)
{
    int k = dev_locate_index(t.delta, t.nb_d, delta);
    int l = dev_locate_index(t.gamma, t.nb_g, gamma);
    double tt, u;
    if (k < 0) { tt = 0.; k = 0; }
    else if (k > t.nb_d - 2) { tt = 1.; k = t.nb_d - 2; }
    else tt = (delta - t.delta[k]) / (t.delta[k + 1] - t.delta[k]);
    if (l < 0) { l = 0; u = 0.; }
    else if (l > t.nb_g - 2) { l = t.nb_g - 2; u = 1.; }
    else u = (gamma - t.gamma[l]) / (t.gamma[l + 1] - t.gamma[l]);
    const double* p = t.p + (size_t)k * t.nb_g + l;
    double e = p[0] * (1. - tt) * (1. - u) + p[t.nb_g] * tt * (1. - u) + p[1] * (1. - tt) * u + p[t.nb_g + 1] * u * tt;
    if (e > 1.) e = 1.;
    else if (e < 0.) e = 0.;
    return e;
}

// lvg_line_overlap_data::get_esc_func, lvg_method_functions.cpp:324-392: 16-corner multilinear interpolation in
// (log10 delta, dx, gamma ratio, gamma); same clamps; same order of the 16 terms.  Split into the four axis searches
// (OvlLoc) and the table evaluation so that two tables on identical grids share one search.
struct OvlLoc { int m, n, k, l; double y, p, t, u; };

__device__ __forceinline__ OvlLoc dev_overlap_locate(const DevOverlap& o, double gamma, double log_delta, double gamma_ratio,
                                                     double delta_x)
{
    OvlLoc q;
    q.m = dev_locate_index(o.log_delta, o.nb_d, log_delta);
    q.l = dev_locate_index(o.gamma, o.nb_g, gamma);
    q.k = dev_locate_index(o.gratio, o.nb_gr, gamma_ratio);
    q.n = dev_locate_index(o.dx, o.nb_dx, delta_x);
    q.y = q.u = q.t = q.p = 0.;
    if (q.m < 0) q.m = 0;
    else if (q.m > o.nb_d - 2) { q.m = o.nb_d - 2; q.y = 1.; }
    else q.y = (log_delta - o.log_delta[q.m]) / (o.log_delta[q.m + 1] - o.log_delta[q.m]);
    if (q.n < 0) q.n = 0;
    else if (q.n > o.nb_dx - 2) { q.p = 1.; q.n = o.nb_dx - 2; }
    else q.p = (delta_x - o.dx[q.n]) / (o.dx[q.n + 1] - o.dx[q.n]);
    if (q.l < 0) q.l = 0;
    else if (q.l > o.nb_g - 2) { q.l = o.nb_g - 2; q.u = 1.; }
    else q.u = (gamma - o.gamma[q.l]) / (o.gamma[q.l + 1] - o.gamma[q.l]);
    if (q.k < 0) q.k = 0;
    else if (q.k > o.nb_gr - 2) { q.t = 1.; q.k = o.nb_gr - 2; }
    else q.t = (gamma_ratio - o.gratio[q.k]) / (o.gratio[q.k + 1] - o.gratio[q.k]);
    return q;
}

__device__ __forceinline__ double dev_overlap_eval(const DevOverlap& o, const OvlLoc& q)
{
    const size_t W = (size_t)o.nb_g * o.nb_gr;
    const double* __restrict__ tab = o.p;
    double v[16];
#pragma unroll
    for (int a = 0; a < 2; a++)
#pragma unroll
        for (int b = 0; b < 2; b++)
#pragma unroll
            for (int c = 0; c < 2; c++)
#pragma unroll
                for (int e = 0; e < 2; e++)
                    v[((a * 2 + b) * 2 + c) * 2 + e] = tab[(size_t)((q.m + a) * o.nb_dx + q.n + b) * W + (size_t)(q.k + c) * o.nb_g + q.l + e];
    double escf = 0.;
#pragma unroll
    for (int a = 0; a < 2; a++) {          // delta corner (y)
#pragma unroll
        for (int b = 0; b < 2; b++) {      // dx corner (p)
#pragma unroll
            for (int c = 0; c < 2; c++) {  // gamma-ratio corner (t)
#pragma unroll
                for (int e = 0; e < 2; e++)  // gamma corner (u)
                    escf += v[((a * 2 + b) * 2 + c) * 2 + e] * (e ? q.u : 1. - q.u) * (c ? q.t : 1. - q.t) * (b ? q.p : 1. - q.p) * (a ? q.y : 1. - q.y);
            }
        }
    }
    if (escf > 1.) escf = 1.;
    else if (escf < 0.) escf = 0.;
    return escf;
}

__device__ __forceinline__ double dev_overlap_esc_func(const DevOverlap& o, double gamma, double delta, double gamma_ratio,
                                                       double delta_x)
{
    return dev_overlap_eval(o, dev_overlap_locate(o, gamma, log10(delta), gamma_ratio, delta_x));
}

// Per-layer scalars of one coupling evaluation
struct LayerCtx {
    double mol_conc, vel_width, vel_grad; // vel_grad signed
    const double* dust_conc;              // [ncomp]
};

__device__ __forceinline__ LayerCtx dev_layer_ctx(const DevModel& d, int lay)
{
    LayerCtx c;
    c.mol_conc = d.mol_conc[lay];
    c.vel_width = d.vel_width[lay];
    c.vel_grad = d.vel_grad[lay];
    c.dust_conc = d.dust_conc + (size_t)lay * d.ncomp;
    return c;
}

// dust_model::absorption(energy, conc), dust_model.cpp:834-841, with dust_component::absorption(energy)
// (:473-490) evaluated once per line on the host (it depends on the line energy only) -> sigma[ncomp].
__device__ __forceinline__ double dev_dust_opacity(const DevModel& d, const LayerCtx& c, const double* __restrict__ sigma)
{
    double a = 0.;
    for (int k = 0; k < d.ncomp; k++) a += sigma[k] * c.dust_conc[k];
    return a;
}

// iteration_scheme_lvg::intensity_calc, iteration_lvg.cpp:163-185
// `exact`: the net absorption as the difference of two ROUNDED products, operation by operation as the reference's expression
// evaluates without FMA hardware (iteration_lvg.cpp:171-172).  The two forms differ by the rounding of one product, which is
// the whole answer for a line AT the inversion threshold (absorption = stimulated emission: J = S ep reaches 1e8 and moves by
// 1e-5 relative; seen at BASELINE size in lim_luminosity_lvg).  The one-shot evaluations (rt_intensity, rt_lim_luminosity)
// use the exact form.  The ITERATION keeps the contracted form on purpose: the difference of rounded products can cancel to
// exactly 0, leaving line_opacity = MIN_LINE_OPACITY = 1e-99 and J ~ 1e80 -- the un-contracted iteration walked two of the
// rounding-sensitive layers of the BASELINE slabs into such states (negative populations flagged as converged; a next step
// with eq_error 1e53); the fused multiply-subtract leaves the product's rounding residual instead and stays finite.
__device__ __forceinline__ double dev_intensity_single(const DevModel& d, const DevLvg& lvg, const LayerCtx& c, double energy,
                                                       double Aul, double Alu, double n_u, double n_l, const double* sigma,
                                                       bool exact = false)
{
    const double cc = c.mol_conc / (d.prm.eight_pi * c.vel_width * energy * energy * energy);
    const double line_emiss = cc * Aul * n_u;
    double line_opacity = exact ? __dadd_rn(__dsub_rn(__dmul_rn(__dmul_rn(cc, Alu), n_l), line_emiss), d.prm.min_line_opacity)
                                : cc * Alu * n_l - line_emiss + d.prm.min_line_opacity;
    if (line_opacity < 0.) line_opacity *= d.prm.inv_trans_factor;
    const double dust_opacity = dev_dust_opacity(d, c, sigma);
    const double avg = fabs(c.vel_grad);
    const double gamma = avg / (c.vel_width * line_opacity);
    const double delta = avg / (c.vel_width * dust_opacity);
    const double ep1 = dev_esc_func(lvg, gamma, delta);
    return line_emiss / line_opacity * ep1;
}
__device__ __forceinline__ double dev_intensity_single(const DevModel& d, const LayerCtx& c, double energy, double Aul,
                                                       double Alu, double n_u, double n_l, const double* sigma, bool exact = false)
{
    return dev_intensity_single(d, d.lvg, c, energy, Aul, Alu, n_u, n_l, sigma, exact);
}

// iteration_scheme_line_overlap::intensity_calc(u1,l1,u2,l2,...), iteration_lvg.cpp:428-501.
// dxe = (E_u1 - E_l1 - E_u2 + E_l2) * c / E1 precomputed on the host in the reference's operation order.
__device__ __forceinline__ void dev_intensity_pair(const DevModel& d, const LayerCtx& c, double energy1, double dx_num,
                                                   double A1ul, double A1lu, double A2ul, double A2lu, double n_u1,
                                                   double n_l1, double n_u2, double n_l2, const double* sigma1,
                                                   double& intens1, double& intens2, bool exact = false)
{
    const double max_dx = d.prm.max_dx;
    double cc = c.mol_conc / (d.prm.eight_pi * c.vel_width * energy1 * energy1 * energy1);
    const double line_emiss1 = cc * A1ul * n_u1;
    double line_opacity1 = exact ? __dadd_rn(__dmul_rn(cc, __dsub_rn(__dmul_rn(A1lu, n_l1), __dmul_rn(A1ul, n_u1))), d.prm.min_line_opacity)
                                 : cc * (A1lu * n_l1 - A1ul * n_u1) + d.prm.min_line_opacity;
    const double line_emiss2 = cc * A2ul * n_u2;
    double line_opacity2 = exact ? __dadd_rn(__dmul_rn(cc, __dsub_rn(__dmul_rn(A2lu, n_l2), __dmul_rn(A2ul, n_u2))), d.prm.min_line_opacity)
                                 : cc * (A2lu * n_l2 - A2ul * n_u2) + d.prm.min_line_opacity;
    if (line_opacity1 < 0.) line_opacity1 *= d.prm.inv_trans_factor;
    if (line_opacity2 < 0.) line_opacity2 *= d.prm.inv_trans_factor;
    const double avg = fabs(c.vel_grad);
    const double gamma1 = avg / (c.vel_width * line_opacity1);
    const double gamma2 = avg / (c.vel_width * line_opacity2);
    const double dust_opacity = dev_dust_opacity(d, c, sigma1);
    const double delta = avg / (c.vel_width * dust_opacity);
    // dx = (E1 - E2) * SPEED_OF_LIGHT / (energy * vel_width)
    double dx = dx_num / (energy1 * c.vel_width);
    if (c.vel_grad < 0.) dx *= -1.;
    double ep1 = 0., ep2 = 0., ep01 = 0., ep02 = 0.;
    const double adx = fabs(dx);
    if (adx < max_dx) {
        ep1 = dev_overlap_esc_func(d.ov[0], gamma1, delta, gamma2 / gamma1, dx);
        ep2 = dev_overlap_esc_func(d.ov[0], gamma2, delta, gamma1 / gamma2, -dx);
    }
    if (adx > max_dx - 0.5) {
        ep01 = dev_esc_func(d.lvg, gamma1, delta);
        ep02 = dev_esc_func(d.lvg, gamma2, delta);
    }
    if (adx > max_dx) { ep1 = ep01; ep2 = ep02; }
    else if (adx > max_dx - 0.5) {
        cc = 2. * (max_dx - adx);
        ep1 = ep01 * (1. - cc) + ep1 * cc;
        ep2 = ep02 * (1. - cc) + ep2 * cc;
    }
    intens1 = line_emiss1 / line_opacity1 * ep1;
    intens2 = line_emiss2 / line_opacity2 * ep2;
    if (adx < max_dx) {
        ep1 = dev_overlap_esc_func(d.ov[1], gamma1, delta, gamma2 / gamma1, dx);
        ep2 = dev_overlap_esc_func(d.ov[1], gamma2, delta, gamma1 / gamma2, -dx);
        if (adx > max_dx - 0.5) {
            cc = 2. * (max_dx - adx);
            ep1 *= cc;
            ep2 *= cc;
        }
        intens1 += line_emiss2 / line_opacity2 * ep1;
        intens2 += line_emiss1 / line_opacity1 * ep2;
    }
}

// ---- the pair evaluation split in two independent halves (fused_kernels.cu runs them on two threads) --------------------
// Half X (0: line 1, 1: line 2) looks up everything that depends on (gamma_X, delta, gamma_Y / gamma_X, +-dx):
// e[0] = overlap table 1, e[1] = overlap table 2, e[2] = single-line table; dev_pair_combine then repeats the blending of
// dev_intensity_pair operation by operation, so the two routes give bit-identical intensities.
struct PairCommon { double line_emiss1, line_opacity1, line_emiss2, line_opacity2, gamma1, gamma2, delta, dx, adx; };

__device__ __forceinline__ PairCommon dev_pair_common(const DevModel& d, const LayerCtx& c, double energy1, double dx_num, double A1ul,
                                                      double A1lu, double A2ul, double A2lu, double n_u1, double n_l1, double n_u2,
                                                      double n_l2, const double* sigma1)
{
    PairCommon q;
    const double cc = c.mol_conc / (d.prm.eight_pi * c.vel_width * energy1 * energy1 * energy1);
    q.line_emiss1 = cc * A1ul * n_u1;
    q.line_opacity1 = cc * (A1lu * n_l1 - A1ul * n_u1) + d.prm.min_line_opacity;
    q.line_emiss2 = cc * A2ul * n_u2;
    q.line_opacity2 = cc * (A2lu * n_l2 - A2ul * n_u2) + d.prm.min_line_opacity;
    if (q.line_opacity1 < 0.) q.line_opacity1 *= d.prm.inv_trans_factor;
    if (q.line_opacity2 < 0.) q.line_opacity2 *= d.prm.inv_trans_factor;
    const double avg = fabs(c.vel_grad);
    q.gamma1 = avg / (c.vel_width * q.line_opacity1);
    q.gamma2 = avg / (c.vel_width * q.line_opacity2);
    const double dust_opacity = dev_dust_opacity(d, c, sigma1);
    q.delta = avg / (c.vel_width * dust_opacity);
    q.dx = dx_num / (energy1 * c.vel_width);
    if (c.vel_grad < 0.) q.dx *= -1.;
    q.adx = fabs(q.dx);
    return q;
}

__device__ __forceinline__ void dev_pair_half(const DevModel& d, const DevLvg& lvg, const DevOverlap& ov0, const DevOverlap& ov1,
                                              bool same_axes, const PairCommon& q, int half, double (&e)[3])
{
    const double max_dx = d.prm.max_dx;
    const double gx = half ? q.gamma2 : q.gamma1, gy = half ? q.gamma1 : q.gamma2, dxx = half ? -q.dx : q.dx;
    e[0] = e[1] = e[2] = 0.;
    if (q.adx < max_dx) {
        const double ld = log10(q.delta);
        const OvlLoc l0 = dev_overlap_locate(ov0, gx, ld, gy / gx, dxx);
        e[0] = dev_overlap_eval(ov0, l0);
        e[1] = dev_overlap_eval(ov1, same_axes ? l0 : dev_overlap_locate(ov1, gx, ld, gy / gx, dxx));
    }
    if (q.adx > max_dx - 0.5) e[2] = dev_esc_func(lvg, gx, q.delta);
}

__device__ __forceinline__ void dev_pair_combine(const DevModel& d, const PairCommon& q, const double (&a)[3], const double (&b)[3],
                                                 double& intens1, double& intens2)
{
    const double max_dx = d.prm.max_dx, adx = q.adx;
    double ep1 = a[0], ep2 = b[0];
    const double ep01 = a[2], ep02 = b[2];
    double cc;
    if (adx > max_dx) { ep1 = ep01; ep2 = ep02; }
    else if (adx > max_dx - 0.5) {
        cc = 2. * (max_dx - adx);
        ep1 = ep01 * (1. - cc) + ep1 * cc;
        ep2 = ep02 * (1. - cc) + ep2 * cc;
    }
    intens1 = q.line_emiss1 / q.line_opacity1 * ep1;
    intens2 = q.line_emiss2 / q.line_opacity2 * ep2;
    if (adx < max_dx) {
        ep1 = a[1]; ep2 = b[1];
        if (adx > max_dx - 0.5) {
            cc = 2. * (max_dx - adx);
            ep1 *= cc;
            ep2 *= cc;
        }
        intens1 += q.line_emiss2 / q.line_opacity2 * ep1;
        intens2 += q.line_emiss1 / q.line_opacity1 * ep2;
    }
}

