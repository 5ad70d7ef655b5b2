"""Seeded synthetic problems of the shapes named in BASELINE.json `configs`.

No spectroscopic, collisional, dust or shock-profile data ship with the reference, so
every molecule here is synthetic but *physically shaped*: rigid-rotor-like ladders, A ~ nu^3, collision rates
~1e-11..1e-10 cm3/s with detailed balance applied by the solver, cloud layers spanning T = 50..300 K and
n_H = 1e5..1e8 cm-3, velocity gradients with a sign change, one silicate-like dust component whose grid starts
above the lowest line energies (so the power-law branch of dust_component::absorption is exercised).

Tables follow the reference's conventions exactly: collision tables are packed i*(i-1)/2 + j with coeff[.,0]=0
at T=0 (e.g. coll_rates_oh.cpp:58); the LVG table is p[delta][gamma] (lvg_method_functions.cpp:37-56).
"""
from __future__ import annotations

import numpy as np

from .abi import (CollTable, DustComp, OverlapTable, Problem, RT_RULE_CH3OH, RT_RULE_GENERIC, RT_RULE_H2CO,
                  RT_RULE_H2O, RT_RULE_NH3, RT_RULE_OH_HF, RT_SCHEME_LINE_OVERLAP, RT_SCHEME_LVG, RT_START_BOUNDARY, RT_START_LTE)

AMU = 1.660539040e-24


# ---------------------------------------------------------------------------------------------------------
# LVG tables
# ---------------------------------------------------------------------------------------------------------
def _beta_slab(gamma: np.ndarray) -> np.ndarray:
    """Two-sided LVG escape probability of a plane-parallel flow (Hummer & Rybicki 1985; the integrand is the
    reference's lvg_func_g with td = 0, lvg_method_functions.h:10-16): gamma * int_0^1 mu^2 (1-exp(-1/(gamma mu^2)))."""
    x, w = np.polynomial.legendre.leggauss(96)
    mu = 0.5 * (x + 1.0)
    w = 0.5 * w
    g = gamma[..., None]
    val = gamma * np.sum(w * mu * mu * (-np.expm1(-1.0 / (g * mu * mu))), axis=-1)
    return np.clip(val, 0.0, 1.0)


def make_lvg_table(nb_d: int = 61, nb_g: int = 81):
    """v(delta, gamma) = (1 - beta(gamma)) / (1 + gamma/delta): fraction of line photons that is neither lost
    through the velocity gradient nor absorbed by dust in one flight.  One-sided escape p = 0.5 (1 - v)."""
    gamma = np.logspace(-8, 5, nb_g)
    delta = np.logspace(-3, 13, nb_d)
    beta = _beta_slab(gamma)
    p = (1.0 - beta)[None, :] / (1.0 + gamma[None, :] / delta[:, None])
    return delta, gamma, np.ascontiguousarray(np.clip(p, 0.0, 1.0))


def make_overlap_tables(nb_d: int = 9, nb_dx: int = 17, nb_gr: int = 9, nb_g: int = 25):
    """Synthetic p1/p2 tables of lvg_line_overlap_data (lvg_method_functions.h:53-69).  p1: probability that a
    photon of line 1 is re-absorbed in line 1 given a second line at offset dx with opacity ratio g2/g1;
    p2: probability it is absorbed in line 2.  Smooth, in [0,1], asymmetric in dx (the flow carries photons
    only one way in frequency), and p1 -> the single-line value, p2 -> 0 as |dx| grows."""
    gamma = np.logspace(-6, 4, nb_g)
    delta = np.logspace(-2, 12, nb_d)
    dx = np.linspace(-4.0, 4.0, nb_dx)
    gratio = np.logspace(-3, 3, nb_gr)
    beta = _beta_slab(gamma)
    single = (1.0 - beta)[None, :] / (1.0 + gamma[None, :] / delta[:, None])  # [d, g]
    p1 = np.empty((nb_d * nb_dx, nb_gr * nb_g))
    p2 = np.empty_like(p1)
    for l in range(nb_d):
        for k in range(nb_dx):
            # overlap weight: photons of line 1 are red-shifted into line 2 only when dx > 0
            wgt = np.exp(-0.5 * dx[k] ** 2 / 1.5 ** 2) * (0.35 + 0.65 / (1.0 + np.exp(-2.0 * dx[k])))
            for i in range(nb_gr):
                # opacity of line 2 relative to line 1 is 1/gratio
                share = (1.0 / gratio[i]) / (1.0 + 1.0 / gratio[i])
                s = single[l]
                p1[l * nb_dx + k, i * nb_g:(i + 1) * nb_g] = s * (1.0 - wgt * share * 0.8)
                p2[l * nb_dx + k, i * nb_g:(i + 1) * nb_g] = s * wgt * share * 0.8
    mk = lambda p: OverlapTable(delta.copy(), dx.copy(), gratio.copy(), gamma.copy(), np.ascontiguousarray(np.clip(p, 0, 1)))
    return mk(p1), mk(p2)


# ---------------------------------------------------------------------------------------------------------
# dust, cloud
# ---------------------------------------------------------------------------------------------------------
def make_dust(rng: np.random.Generator, ncomp: int = 1, nb_ph_en: int = 120):
    comps = []
    for c in range(ncomp):
        # wavelengths descending in the file -> energies ascending; grid starts at 5 cm-1
        ph_en = np.logspace(np.log10(5.0), 5.0, nb_ph_en) * (1.0 + 0.01 * c)
        area = 3.0e-10 * (1.0 + c)
        q = np.minimum(1.0, (ph_en / 3000.0) ** 1.8) * (1.0 + 0.3 * np.exp(-0.5 * ((np.log(ph_en) - np.log(1000.0)) / 0.25) ** 2))
        q *= np.exp(0.03 * rng.standard_normal(nb_ph_en))
        comps.append(DustComp(ph_en, area * q, 2.0 - 0.2 * c))
    return comps


def make_cloud(rng: np.random.Generator, L: int, ncomp: int, t_lo=50.0, t_hi=300.0, nh_lo=1e5, nh_hi=1e8,
               x_mol=1e-6, velg_scale=1e-10, velg_swing=1.5):
    s = np.linspace(0.0, 1.0, L) if L > 1 else np.array([0.5])
    # hot, compressed shock layer in the middle of the slab; cool wings
    bump = np.exp(-0.5 * ((s - 0.45) / 0.22) ** 2)
    temp_n = t_lo + (t_hi - t_lo) * bump * (1.0 + 0.02 * rng.standard_normal(L))
    nh = nh_lo * (nh_hi / nh_lo) ** (0.15 + 0.85 * np.sqrt(bump)) * np.exp(0.05 * rng.standard_normal(L))
    h2 = 0.5 * nh * 0.98
    opr = 3.0
    lay = dict(
        temp_n=temp_n,
        temp_el=temp_n * (1.0 + 0.3 * bump),
        el_conc=1e-7 * nh,
        h_conc=0.02 * nh,
        ph2_conc=h2 / (1.0 + opr),
        oh2_conc=h2 * opr / (1.0 + opr),
        he_conc=0.09 * nh,
        mol_conc=x_mol * nh * (0.3 + 0.7 * bump),
        vel_turb=np.full(L, 3.0e4),
    )
    # velocity gradient: compressive then expanding, clamped like cloud_data.cpp:370-376 (MIN_VELOCITY_GRADIENT)
    vg = velg_scale * (10.0 ** (velg_swing * np.cos(3.0 * np.pi * s))) * np.sign(0.62 - s + 1e-9)
    vg = np.where(np.abs(vg) < 3e-14, np.sign(vg) * 3e-14, vg)
    lay["velg_n"] = vg
    dz = np.full(L, 2.0e15 / max(L, 1)) * (1.0 + 0.1 * rng.random(L))
    lay["dz"] = dz
    lay["vel_n"] = 2.0e6 * (1.0 - 0.8 * s)
    lay["dust_conc"] = np.ascontiguousarray((1e-12 * nh)[:, None] * (1.0 / (1.0 + np.arange(ncomp)))[None, :])
    lay["height"] = float(np.sum(dz))
    return {k: (np.ascontiguousarray(v, dtype=np.float64) if k != "height" else v) for k, v in lay.items()}


def _coll_table(rng, nb_lev, tgrid, base, tpow=0.3, sigma=0.5):
    """coeff[i*(i-1)/2+j][t] = base[i,j] * lognormal * (T/100)^tpow, 0 at T = 0."""
    iu, ju = np.tril_indices(nb_lev, -1)  # row-major order of (i>j) pairs == packed index order
    b = base[iu, ju] * np.exp(sigma * rng.standard_normal(iu.shape[0]))
    tt = np.where(tgrid > 0, (np.maximum(tgrid, 1e-30) / 100.0) ** tpow, 0.0)
    wiggle = 1.0 + 0.1 * rng.standard_normal((iu.shape[0], tgrid.shape[0]))
    return CollTable(nb_lev, tgrid.astype(np.float64), np.ascontiguousarray(b[:, None] * tt[None, :] * np.abs(wiggle)))


# ---------------------------------------------------------------------------------------------------------
# molecules
# ---------------------------------------------------------------------------------------------------------
def _einst_from_lines(N, g, energy, lines):
    """lines: list of (u, l, A_ul).  arr[u][l] = A, arr[l][u] = A g_u/g_l (spectroscopy.h:176-178)."""
    e = np.zeros((N, N))
    for u, l, a in lines:
        assert u > l and energy[u] > energy[l]
        e[u, l] = a
        e[l, u] = a * g[u] / g[l]
    return e


def make_oh_hf(rng: np.random.Generator, nb_rot: int = 14):
    """OH-like molecule with Lambda doubling and hyperfine splitting: N = 4*nb_rot levels, levels (2k, 2k+1) are
    the two F components of one (rotational, parity) state -- the pairing iteration_scheme_line_overlap
    assumes (iteration_lvg.cpp:325-335)."""
    e_rot = np.array([0.0, 83.7, 126.3, 187.5, 201.9, 289.0, 355.1, 429.5, 544.8, 608.0, 782.0, 845.0, 1030.0, 1100.0,
                      1290.0, 1380.0, 1600.0, 1700.0])[:nb_rot]
    j_rot = np.array([1.5, 2.5, 0.5, 1.5, 3.5, 2.5, 4.5, 3.5, 5.5, 4.5, 6.5, 5.5, 7.5, 6.5, 8.5, 7.5, 9.5, 8.5])[:nb_rot]
    ladder = np.array([0, 0, 1, 1, 0, 1, 0, 1, 0, 1, 0, 1, 0, 1, 0, 1, 0, 1])[:nb_rot]
    energy, g, jq, vq, meta = [], [], [], [], []
    for r in range(nb_rot):
        lam = 0.055 * (1.0 + 0.9 * r) * (1.0 + 0.05 * rng.standard_normal())
        for par in (0, 1):
            hf = 0.0018 * (1.0 + 0.4 * rng.random()) * (1.0 + 0.3 * par)
            for fi, F in enumerate((j_rot[r] - 0.5, j_rot[r] + 0.5)):
                energy.append(e_rot[r] + par * lam + fi * hf)
                g.append(int(round(2 * F + 1)))
                jq.append(j_rot[r]); vq.append(0)
                meta.append((r, par, F))
    energy = np.array(energy); order = np.argsort(energy, kind="stable")
    assert np.all(order == np.arange(len(energy))), "levels must already be ascending"
    N = len(energy)
    g = np.array(g, dtype=np.int32)
    lines = []
    for u in range(N):
        ru, pu, Fu = meta[u]
        for l in range(u):
            rl, pl, Fl = meta[l]
            if pu == pl or abs(Fu - Fl) > 1.0 + 1e-9:
                continue
            de = energy[u] - energy[l]
            if ru == rl:
                # Lambda-doublet (18 cm-like) lines: magnetic-dipole-weak, A ~ 1e-11..1e-10
                a = 7.0e-11 * (de / 0.055) ** 3 * (0.1 if Fu != Fl else 1.0)
            elif abs(j_rot[ru] - j_rot[rl]) <= 1.0 + 1e-9:
                s = 1.0 if ladder[ru] == ladder[rl] else 0.08
                a = 2.0e-7 * de ** 3 * s * (0.15 if abs(Fu - Fl) > 0.5 and abs(j_rot[ru] - j_rot[rl]) < 0.5 else 1.0)
                a *= np.exp(0.3 * rng.standard_normal())
            else:
                continue
            lines.append((u, l, float(a)))
    einst = _einst_from_lines(N, g, energy, lines)
    # collision base rates: fall with |dJ| and favour parity-conserving, hyperfine-conserving collisions
    base = np.zeros((N, N))
    for u in range(N):
        ru, pu, Fu = meta[u]
        for l in range(u):
            rl, pl, Fl = meta[l]
            dj = abs(j_rot[ru] - j_rot[rl])
            base[u, l] = 6e-11 * np.exp(-dj / 1.5) * (1.0 if pu == pl else 0.4) * (1.0 if abs((Fu - j_rot[ru]) - (Fl - j_rot[rl])) < 0.1 else 0.3)
    return dict(energy=energy, g=g, v=np.array(vq, dtype=np.int32), j=np.array(jq), einst=einst, base=base, mass=17.0 * AMU)


def make_ch3oh(rng: np.random.Generator, n_vt: int = 3, n_k: int = 16, n_j: int = 16):
    """CH3OH-A-like symmetric-top ladders: n_vt torsional states x n_k K-ladders x n_j J values (default 768)."""
    lev = []
    for vt in range(n_vt):
        for k in range(n_k):
            for jj in range(k, k + n_j):
                e = 0.807 * jj * (jj + 1) + 3.45 * k * k * (1.0 - 0.12 * vt) + vt * (208.0 + 45.0 * vt)
                lev.append((e + 1e-3 * rng.random(), vt, jj, k))
    lev.sort(key=lambda t: t[0])
    energy = np.array([t[0] for t in lev])
    energy -= energy[0]
    vq = np.array([t[1] for t in lev], dtype=np.int32)
    jq = np.array([float(t[2]) for t in lev])
    kq = np.array([t[3] for t in lev])
    g = (4 * (2 * jq + 1)).astype(np.int32)  # (2I+1)(2J+1), I = 3/2 for A species
    N = len(lev)
    lines = []
    for u in range(N):
        for l in range(u):
            dj, dk, dv = abs(jq[u] - jq[l]), abs(kq[u] - kq[l]), abs(int(vq[u]) - int(vq[l]))
            if dj > 1 or dk > 1 or dv > 1 or (dj == 0 and dk == 0):
                continue
            if dv == 1 and not (dk == 1):
                continue
            de = energy[u] - energy[l]
            if de < 0.02:
                continue
            s = (1.0 if dk == 0 else 0.35) * (1.0 if dv == 0 else 0.02)
            a = min(3.0e-7 * de ** 3 * s * np.exp(0.5 * rng.standard_normal()), 5.0)
            lines.append((u, l, float(a)))
    einst = _einst_from_lines(N, g, energy, lines)
    dj = np.abs(jq[:, None] - jq[None, :]); dk = np.abs(kq[:, None] - kq[None, :])
    dv = np.abs(vq[:, None] - vq[None, :])
    base = 1.2e-10 * np.exp(-dj / 2.5 - dk / 1.5) * np.where(dv == 0, 1.0, 0.01 ** dv)
    return dict(energy=energy, g=g, v=vq, j=jq, einst=einst, base=base, mass=32.0 * AMU)


def make_h2o(rng: np.random.Generator, N: int = 411):
    """o-H2O-like asymmetric top: J_tau ladders in up to 5 vibrational states, sorted by energy."""
    lev = []
    vib = [0.0, 1594.7, 3151.6, 3657.1, 3755.9]
    for vi, ev in enumerate(vib):
        for jj in range(0, 22):
            for tau in range(-jj, jj + 1, 2):  # ortho-like half of the 2J+1 tau states
                e = ev + 14.5 * jj * (jj + 1) * (0.62 + 0.38 * (tau + jj) / max(1, 2 * jj)) + 1e-3 * rng.random()
                lev.append((e, vi, jj, tau))
    lev.sort(key=lambda t: t[0])
    lev = lev[:N]
    energy = np.array([t[0] for t in lev]); energy -= energy[0]
    vq = np.array([t[1] for t in lev], dtype=np.int32)
    jq = np.array([float(t[2]) for t in lev]); tq = np.array([t[3] for t in lev])
    g = (3 * (2 * jq + 1)).astype(np.int32)
    lines = []
    for u in range(N):
        for l in range(u):
            dj, dt, dv = abs(jq[u] - jq[l]), abs(tq[u] - tq[l]), abs(int(vq[u]) - int(vq[l]))
            if dj > 1 or dt > 4 or dv > 1 or (dt % 4 == 0 and dj == 0 and dv == 0):
                continue
            de = energy[u] - energy[l]
            if de < 0.05:
                continue
            s = (1.0 if dt <= 2 else 0.1) * (1.0 if dv == 0 else 0.05)
            if rng.random() > (0.9 if dt <= 2 else 0.35):
                continue
            lines.append((u, l, float(min(1.0e-6 * de ** 3 * s * np.exp(0.6 * rng.standard_normal()), 50.0))))
    einst = _einst_from_lines(N, g, energy, lines)
    dj = np.abs(jq[:, None] - jq[None, :]); dt = np.abs(tq[:, None] - tq[None, :]); dv = np.abs(vq[:, None] - vq[None, :])
    base = 1.5e-10 * np.exp(-dj / 2.0 - dt / 6.0) * np.where(dv == 0, 1.0, 0.003)
    return dict(energy=energy, g=g, v=vq, j=jq, einst=einst, base=base, mass=18.0 * AMU)


# ---------------------------------------------------------------------------------------------------------
# the five BASELINE.json configs (+ reduced sizes for fast parity tests)
# ---------------------------------------------------------------------------------------------------------
def _finish(name, mol_name, rule, scheme, mol, tables, nb_neutral, rng, L, ncomp, accel, max_iter, start, cloud_kw=None,
            overlap=False):
    delta, gamma, p = make_lvg_table()
    dust = make_dust(rng, ncomp)
    lay = make_cloud(rng, L, ncomp, **(cloud_kw or {}))
    o1 = o2 = None
    if overlap:
        o1, o2 = make_overlap_tables()
    return Problem(name=name, mol_name=mol_name, mol_mass=mol["mass"], species_rule=rule, scheme=scheme,
                   energy=mol["energy"], g=mol["g"], v=mol["v"], j=mol["j"], einst=mol["einst"], tables=tables,
                   nb_neutral_tables=nb_neutral, lvg_delta=delta, lvg_gamma=gamma, lvg_p=p, dust=dust, layers=lay,
                   overlap1=o1, overlap2=o2, accel=accel, max_iter=max_iter, start_mode=start)


def config_oh(L: int = 100, overlap: bool = False, seed: int = 20240001, nb_rot: int = 14) -> Problem:
    """Configs 1 (no overlap, LTE start) and 3 (hyperfine line overlap, boundary start): OH N = 4*nb_rot = 56,
    un-accelerated, 15000 iterations max (radiative_transfer.cpp:440,571)."""
    rng = np.random.default_rng(seed + (2 if overlap else 0))
    mol = make_oh_hf(rng, nb_rot)
    N = len(mol["energy"])
    t_he = np.concatenate([[0.0], np.linspace(5.0, 350.0, 19)])
    t_h2 = np.concatenate([[0.0], np.linspace(10.0, 300.0, 11)])
    # He data cover fewer levels than the diagram: exercises the `up_lev.nb < nb_lev` masks (coll_rates_oh.cpp:396-403)
    tabs = [_coll_table(rng, max(4, N - 16), t_he, mol["base"] * 0.6),
            _coll_table(rng, max(4, N - 8), t_h2, mol["base"] * 1.0),
            _coll_table(rng, max(4, N - 8), t_h2, mol["base"] * 1.4)]
    return _finish("oh_overlap" if overlap else "oh_nooverlap", "OH", RT_RULE_OH_HF,
                   RT_SCHEME_LINE_OVERLAP if overlap else RT_SCHEME_LVG, mol, tabs, 3, rng, L, 1, False, 15000,
                   RT_START_BOUNDARY if overlap else RT_START_LTE, cloud_kw=dict(x_mol=3e-7), overlap=overlap)


def config_ch3oh(L: int = 200, seed: int = 20240002, n_vt: int = 3, n_k: int = 16, n_j: int = 16,
                 cloud_kw=None) -> Problem:
    """Config 2: CH3OH-A N = n_vt*n_k*n_j = 768, 3 tables (He 41, p-H2 21, o-H2 21 temperatures) mixed by the
    vt / J<=9 rule (coll_rates_ch3oh.cpp:484-533), accelerated, 150 iterations, boundary start."""
    rng = np.random.default_rng(seed)
    mol = make_ch3oh(rng, n_vt, n_k, n_j)
    N = len(mol["energy"])
    t_he = np.concatenate([[0.0], np.linspace(10.0, 400.0, 40)])
    t_h2 = np.concatenate([[0.0], np.linspace(10.0, 200.0, 20)])
    tabs = [_coll_table(rng, N, t_he, mol["base"] * 0.5), _coll_table(rng, N, t_h2, mol["base"]),
            _coll_table(rng, N, t_h2, mol["base"] * 1.3)]
    return _finish("ch3oh_a", "CH3OHa", RT_RULE_CH3OH, RT_SCHEME_LVG, mol, tabs, 3, rng, L, 1, True, 150,
                   RT_START_BOUNDARY, cloud_kw=cloud_kw)


H2O_THICK_CLOUD = dict(x_mol=3e-5, t_hi=1200.0)  # round-1 cloud: lines so thick that the accelerated iteration fails in
# most layers -- kept for the small retry-policy cases (tests: h2o_small), where that failure is the point
H2O_CLOUD = dict(x_mol=3e-7, t_hi=1200.0, velg_scale=2e-8, velg_swing=0.7)


def config_h2o(L: int = 200, N: int = 411, seed: int = 20240004, cloud_kw=None) -> Problem:
    """Config 4: o-H2O, 7 tables incl. electrons with the `nb < 45` rule (coll_rates_h2o.cpp:494-548).
    Default cloud: a 1200 K shock layer with x(H2O) = 3e-7 and |dv/dz| = 4e-9 .. 1e-7 s-1, for which the reference
    converges every layer within the 150 accelerated iterations (mean 36, max 71 at N = 411: the Ng steps do fire) and only
    ~5 % of the layers have an answer that is decided by rounding (tests/full_slab_cases.py; a thicker cloud -- x = 1e-6 --
    has 15 % such layers, on which the Ng extrapolation of the reference itself ends in NaN under a last-bit perturbation) (so the slab measures the batched LU, not the
    15000-iteration retry policy of radiative_transfer.cpp:259-276) while dozens of lines are inverted."""
    rng = np.random.default_rng(seed)
    mol = make_h2o(rng, N)
    n45 = min(N, 45)
    tg = lambda n, hi: np.concatenate([[0.0], np.linspace(20.0, hi, n - 1)])
    b = mol["base"]
    tabs = [_coll_table(rng, n45, tg(11, 1500.0), b * 0.5),       # 0 He
            _coll_table(rng, N, tg(10, 4000.0), b * 0.4),         # 1 He rovibrational (He + 0.2 H)
            _coll_table(rng, n45, tg(9, 1500.0), b * 1.0),        # 2 p-H2
            _coll_table(rng, n45, tg(9, 1500.0), b * 1.5),        # 3 o-H2
            _coll_table(rng, N, tg(12, 4000.0), b * 1.2),         # 4 H2 rovibrational (p + o)
            _coll_table(rng, n45, tg(15, 1500.0), b * 2.0),       # 5 H
            _coll_table(rng, N, tg(12, 8000.0), b * 3.0e3)]       # 6 electrons
    return _finish("h2o_ortho", "oH2O", RT_RULE_H2O, RT_SCHEME_LVG, mol, tabs, 6, rng, L, 1, True, 150, RT_START_BOUNDARY,
                   cloud_kw=dict(H2O_CLOUD if cloud_kw is None else cloud_kw))


def config_nh3(L: int = 50, seed: int = 20240006) -> Problem:
    """NH3-like molecule (N = 42): 4 neutral tables He / p-H2 / o-H2 / H mixed by
    nh3_collisions::get_rate_neutrals (coll_rates_nh3.cpp:273-290): the three hydrogen tables share the level limit of the
    p-H2 table, which the tables below make visible (He covers fewer levels, o-H2 and H more)."""
    rng = np.random.default_rng(seed)
    mol = make_ch3oh(rng, 1, 6, 7)
    N = len(mol["energy"])
    tg = lambda n, hi: np.concatenate([[0.0], np.linspace(15.0, hi, n - 1)])
    b = mol["base"]
    tabs = [_coll_table(rng, N - 10, tg(16, 300.0), b * 0.5), _coll_table(rng, N - 4, tg(12, 200.0), b),
            _coll_table(rng, N, tg(12, 200.0), b * 1.3), _coll_table(rng, N, tg(9, 500.0), b * 0.8)]
    return _finish("nh3_para", "pNH3", RT_RULE_NH3, RT_SCHEME_LVG, mol, tabs, 4, rng, L, 1, True, 150, RT_START_BOUNDARY)


def config_h2co(L: int = 50, seed: int = 20240007) -> Problem:
    """H2CO-like molecule (N = 40): 3 neutral tables He / p-H2 / o-H2, h2co_collisions::get_rate_neutrals
    (coll_rates_h2co.cpp:170-185)."""
    rng = np.random.default_rng(seed)
    mol = make_ch3oh(rng, 1, 5, 8)
    N = len(mol["energy"])
    tg = lambda n, hi: np.concatenate([[0.0], np.linspace(10.0, hi, n - 1)])
    b = mol["base"]
    tabs = [_coll_table(rng, N - 8, tg(14, 300.0), b * 0.5), _coll_table(rng, N - 2, tg(11, 300.0), b),
            _coll_table(rng, N, tg(11, 300.0), b * 1.4)]
    return _finish("h2co_ortho", "oH2CO", RT_RULE_H2CO, RT_SCHEME_LVG, mol, tabs, 3, rng, L, 1, True, 150, RT_START_BOUNDARY)


def sweep_cloud_kw(i_n: int, i_t: int, i_c: int, grid: int = 8) -> dict:
    """Config 5: cloud parameters of one point of the density x temperature x specific-column-density grid
    (n_H2 in logspace(4,8), T in linspace(30,300), n_mol/|dv/dz| through |dv/dz| in logspace(-11,-9))."""
    nh = 10.0 ** np.linspace(4.0, 8.0, grid)[i_n]
    tt = np.linspace(30.0, 300.0, grid)[i_t]
    vg = 10.0 ** np.linspace(-11.0, -9.0, grid)[i_c]
    return dict(nh_lo=nh / 30.0, nh_hi=nh, t_lo=tt / 3.0, t_hi=tt, velg_scale=vg)


def sweep_layers(point: int, grid: int, L: int = 200, ncomp: int = 1, seed: int = 20240005) -> dict:
    """The layer arrays (only) of grid point `point` = (i_n * grid + i_t) * grid + i_c: in a sweep the molecule and its
    tables are uploaded once per GPU and only the cloud changes, so the cloud has its own seed."""
    i_n, r = divmod(int(point), grid * grid)
    i_t, i_c = divmod(r, grid)
    rng = np.random.default_rng(seed + 7919 * int(point))
    return make_cloud(rng, L, ncomp, **sweep_cloud_kw(i_n, i_t, i_c, grid))


def config_sweep_point(i_n: int, i_t: int, i_c: int, grid: int = 8, L: int = 200, seed: int = 20240002, **kw) -> Problem:
    """Config 5: one point of the grid as a complete Problem (same molecule as config 2; only the cloud changes)."""
    p = config_ch3oh(L=L, seed=seed, **kw)
    p.layers = sweep_layers((i_n * grid + i_t) * grid + i_c, grid, L=L, ncomp=len(p.dust))
    return p


def small(name: str, L: int = 6) -> Problem:
    """Reduced-size variants used by the parity tests (seconds on one CPU core)."""
    if name == "oh":
        return config_oh(L=L, overlap=False)
    if name == "oh_overlap":
        return config_oh(L=L, overlap=True)
    if name == "ch3oh_small":
        return config_ch3oh(L=L, n_vt=2, n_k=6, n_j=8)  # N = 96
    if name == "ch3oh_mid":
        return config_ch3oh(L=L, n_vt=3, n_k=8, n_j=10)  # N = 240
    if name == "h2o_small":
        return config_h2o(L=L, N=150, cloud_kw=H2O_THICK_CLOUD)
    if name in ("edge4", "edge9", "edge32", "edge33", "edge62", "edge64", "edge98", "edge128", "edge136", "edge152", "edge161"):
        # level counts at the seams of the solver's size classes (row slots per lane of the fused solver's register LU):
        # 4 and 9 (fewer levels than one column block; odd), 32, 33, 62 and 64 (the ends of the one- and two-slot classes),
        # 98 (first of the four-slot class), 128 and 136 (multiples of 8: the right-hand side is alone in the last column
        # block; 128 fills the four row slots, 136 is in the five-slot class), 152 (above the register LU: LU in shared
        # memory), 161 (just above the fused kernel: blocked multi-kernel LU)
        shape = {"edge4": (1, 1, 4), "edge9": (1, 3, 3), "edge32": (1, 4, 8), "edge33": (1, 3, 11), "edge62": (1, 2, 31),
                 "edge64": (1, 8, 8), "edge98": (2, 7, 7), "edge128": (2, 8, 8), "edge136": (1, 8, 17), "edge152": (1, 8, 19),
                 "edge161": (1, 7, 23)}[name]
        return config_ch3oh(L=L, n_vt=shape[0], n_k=shape[1], n_j=shape[2])
    if name == "h2o_odd":
        return config_h2o(L=L, N=131)  # odd level count in the five-row-slot class of the fused solver's register LU
    if name == "nh3":
        return config_nh3(L=L)
    if name == "h2co":
        return config_h2co(L=L)
    if name in ("generic", "generic_spline"):
        p = config_ch3oh(L=L, n_vt=1, n_k=5, n_j=6)  # N = 30
        p.species_rule = RT_RULE_GENERIC
        p.mol_name = "synthetic"
        if name == "generic_spline":
            # the cubic-spline table class collision_data_cub_spline (coll_rates.cpp:89-132) for two of
            # the three tables, the third stays linear (both classes side by side in one collisional_transitions)
            p.name = "generic_spline"
            p.tables[0].interp = 1
            p.tables[2].interp = 1
        return p
    raise KeyError(name)
