"""The golden-vector cases: what is computed, on which seeded problem.  Shared by tests/golden/make_golden.py (runs
them through the UNMODIFIED reference, oracle/_ref) and by the tests (run them through the C restatement on the
CPU and through the CUDA library on the GPU and compare with the committed reference outputs)."""
from __future__ import annotations

import hashlib
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = {  # name -> (synthetic.small name, layers)
    "generic": ("generic", 3),
    "oh": ("oh", 3),
    "oh_overlap": ("oh_overlap", 3),
    "ch3oh_small": ("ch3oh_small", 3),
    "h2o_small": ("h2o_small", 2),
    "nh3": ("nh3", 3),     # nh3_collisions / h2co_collisions mixing rules
    "h2co": ("h2co", 3),
    "generic_spline": ("generic_spline", 3),  # f-3: collision_data_cub_spline tables (coll_rates.cpp:89-132)
}


def problem(case):
    from radiative_transfer_b200 import synthetic
    name, L = CASES[case]
    return synthetic.small(name, L=L)


def problem_digest(p) -> str:
    """sha256 over every input array: the fixture is only valid for byte-identical inputs."""
    h = hashlib.sha256()
    arrs = [p.energy, p.g, p.v, p.j, p.einst, p.lvg_delta, p.lvg_gamma, p.lvg_p]
    for t in p.tables:
        arrs += [t.tgrid, t.coeff]
        if getattr(t, "interp", 0):
            arrs.append(np.array([t.interp], dtype=np.int64))
    for c in p.dust:
        arrs += [c.ph_en, c.abs_coeff, np.array([c.wvl_exp])]
    for o in (p.overlap1, p.overlap2):
        if o is not None:
            arrs += [o.delta, o.dx, o.gratio, o.gamma, o.p]
    for k in sorted(p.layers):
        arrs.append(np.asarray(p.layers[k], dtype=np.float64))
    for a in arrs:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def probe_points(p, n=64, seed=7):
    """(gamma, delta) probes spanning the LVG table and both out-of-range sides; overlap probes likewise."""
    rng = np.random.default_rng(seed)
    g = 10.0 ** rng.uniform(np.log10(p.lvg_gamma[0]) - 1, np.log10(p.lvg_gamma[-1]) + 1, n)
    d = 10.0 ** rng.uniform(np.log10(p.lvg_delta[0]) - 1, np.log10(p.lvg_delta[-1]) + 1, n)
    g[:4] = [p.lvg_gamma[0], p.lvg_gamma[-1], p.lvg_gamma[3], p.lvg_gamma[0] * 0.5]
    d[:4] = [p.lvg_delta[0], p.lvg_delta[-1], p.lvg_delta[5], p.lvg_delta[-1] * 2.0]
    gr = 10.0 ** rng.uniform(-4, 4, n)
    dx = rng.uniform(-4.5, 4.5, n)
    return g, d, gr, dx


def maser_lines(p, pop, nmax=6):
    """Radiative lines with the strongest population inversion in any layer (or, failing that, the first lines)."""
    N = p.N
    iu, il = np.nonzero(np.tril(p.einst, -1) > 0)
    inv = pop[:, iu] / p.g[iu] - pop[:, il] / p.g[il]
    score = inv.max(axis=0) / (pop[:, iu].max(axis=0) / p.g[iu] + 1e-300)
    order = np.argsort(-score, kind="stable")[:nmax]
    return iu[order].astype(np.int32), il[order].astype(np.int32)


def compute(chk, p):
    """Everything a checker (oracle.loader.Checker) says about problem p.  Returns a dict of arrays."""
    out = {}
    g, d, gr, dx = probe_points(p)
    out["esc"] = chk.esc_func(p, g, d)
    if p.overlap1 is not None:
        out["ov1"] = chk.overlap_esc_func(p, 1, g, d, gr, dx)
        out["ov2"] = chk.overlap_esc_func(p, 2, g, d, gr, dx)
        out["groups"] = chk.line_groups(p)
    rng = np.random.default_rng(11)
    en = np.concatenate([10.0 ** rng.uniform(-1.5, 4.9, 40), [p.dust[0].ph_en[0] * 0.5, p.dust[0].ph_en[-1] * 2.0]])
    out["dust_energy"] = en
    try:
        out["dust"] = chk.dust_absorption(p, 0, en)
    except AssertionError:
        pass
    out.update(project("coll", chk.collision_rates(p)))
    out["boundary"] = chk.boundary(p)
    out["boltzmann"] = chk.boltzmann(p)
    b = out["boundary"]
    M, out["df"] = chk.assemble(p, b)
    out.update(project("M", M))
    out["step_pop"], out["step_eq"] = chk.step(p, b)
    iu, il = np.nonzero(np.tril(p.einst, -1) > 0)
    sel = np.linspace(0, iu.shape[0] - 1, min(64, iu.shape[0])).astype(int)
    out["int_up"], out["int_low"] = iu[sel].astype(np.int32), il[sel].astype(np.int32)
    out["intensity"] = chk.intensity(p, p.L - 1, out["int_up"], out["int_low"], b[p.L - 1])
    pop, st, _ = chk.solve(p)
    out["solve_pop"], out["solve_iters"], out["solve_conv"] = pop, st["iters"].copy(), st["converged"].copy()
    out["solve_rel"] = st["rel_error"].copy()
    # maser post-pass on the converged populations of THIS checker's reference run is not comparable across
    # implementations to 1e-12 (the stop rule leaves 1e-6 differences), so it is evaluated on the boundary start
    # vector scaled to exaggerate inversions: deterministic input, identical for every implementation
    up, low = maser_lines(p, b)
    out["post_up"], out["post_low"] = up, low
    post = chk.maser_postpass(p, b, up, low)
    for k, v in post.items():
        out["post_" + k] = v
    # pump / loss rates, limiting luminosity, saturation depth (lim_luminosity_lvg, calc_saturation_depth) on the same input
    lim = chk.lim_luminosity(p, b, up, low, beaming_factor=1.0, lay_nb_hg=post["scalars"][:, 4])
    for k, v in lim.items():
        out["lim_" + k] = v
    cp, cst, _ = chk.calc_molecular_populations(p)
    out["chain_pop"], out["chain_iters"], out["chain_conv"] = cp, cst["iters"].copy(), cst["converged"].copy()
    return out


def project(name, M):
    """A [L][N][N] matrix is stored whole when small, else as two fixed random projections (every entry enters
    both with an O(1) weight, so any entry-level discrepancy shows up at the same relative size)."""
    L, N, _ = M.shape
    if N <= 56:
        return {name: M}
    rng = np.random.default_rng(3)
    w1 = rng.uniform(0.5, 1.5, N); w2 = rng.uniform(0.5, 1.5, N)
    return {name + "_absrow": np.abs(M) @ w1, name + "_abscol": np.einsum("i,lij->lj", w2, np.abs(M)),
            name + "_row": M @ w1, name + "_diag": np.einsum("lii->li", M).copy()}


def load(case):
    f = os.path.join(GOLDEN_DIR, case + ".npz")
    z = np.load(f)
    return {k: z[k] for k in z.files}
