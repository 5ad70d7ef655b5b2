"""The BASELINE.json slabs at FULL size, every layer: which seeded problem, and the explicit allow-list of layers that are
exempt from the all-layer parity assertion (layer index + reason).  Shared by tests/golden/make_full_slab.py (runs the
UNMODIFIED reference, oracle/_ref, and commits its answer as tests/golden/full_<case>.npz) and tests/test_gpu_full_slab.py
(runs the CUDA path through the C ABI on the same inputs and compares every layer)."""
from __future__ import annotations

import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _synth():
    from radiative_transfer_b200 import synthetic
    return synthetic


# name -> (builder, mode); mode "independent": every layer from the problem's own start (boundary / LTE), the measured
# configuration; mode "chained": calc_molecular_populations itself (warm-start chain + retry, radiative_transfer.cpp:219-289)
CASES = {
    # configs[0]: OH N = 56, L = 100, iteration_scheme_lvg, un-accelerated, LTE start (radiative_transfer.cpp:549-571)
    "oh": (lambda: _synth().config_oh(L=100, overlap=False), "independent"),
    # configs[1]: CH3OH-A N = 768, L = 200, accelerated, boundary start (the headline slab) + a second seed
    "ch3oh": (lambda: _synth().config_ch3oh(L=200), "independent"),
    "ch3oh_seed2": (lambda: _synth().config_ch3oh(L=200, seed=20240012), "independent"),
    # configs[2]: OH with hyperfine line overlap, N = 56, L = 400, un-accelerated
    "oh_overlap": (lambda: _synth().config_oh(L=400, overlap=True), "independent"),
    # configs[3]: o-H2O N = 411, L = 200, and the level count the reference driver uses (N = 150, :901,1035)
    "h2o": (lambda: _synth().config_h2o(L=200), "independent"),
    "h2o150": (lambda: _synth().config_h2o(L=200, N=150), "independent"),
    # the reference's own start mode at full level count: one warm-start chain over 40 layers of the headline molecule
    "ch3oh_chained": (lambda: _synth().config_ch3oh(L=40), "chained"),
}

# Rounding-sensitivity probes run per case by tests/golden/make_full_slab.py (bits: oracle/shim/linear_algebra.h); the
# N = 768 slabs take minutes of CPU per run, so they get three probes, the others all seven.
def sens_probes(case):
    return (1, 2, 5) if case.startswith("ch3oh") else (1, 2, 3, 4, 5, 6, 7)


SENS_PROBES = (1, 2, 5)  # overridden per case by the generator through sens_probes()
# A layer is ROUNDING-SENSITIVE when the reference's own answer moves under the probes -- convergence flag changes, the
# iteration count moves by more than one, the probe run ends in NaN (the Ng step's 4 x 4 system turns singular, as on layer
# 108 of the headline slab), or the populations move by more than SENS_TOL.  On healthy layers the probes move the populations
# by 1e-15 .. 1e-11 (median 1e-11, the size of the GPU-vs-reference differences there); SENS_TOL = 1e-8 is a thousand times
# that median and a hundredth of the parity tolerance.  The probes change the rounding of the solve in three to seven
# specific ways; a blocked, fused, tensor-core solve changes it in every operation, so a layer on which the probes already
# reach 1e-8 is not held to 1e-6.  Measured: o-H2O N = 150 -- the two layers on which the CUDA path is outside 1e-6 (77: 7e-6,
# 86: 4e-5) are the two most sensitive of the slab (2e-5, 5e-5); o-H2O N = 411 -- all nine outliers of the CUDA path are among
# the ten most sensitive layers (five of them end in NaN in the reference under a probe); CH3OH seed 2 -- outliers 6, 18, 26
# have probe deviations 2e-5, 1e-4, 6e-8 against a slab median of 1e-11.
SENS_TOL = 1e-8


def sensitive(gold):
    """Boolean [L]: layers whose reference answer is decided by rounding (see SENS_TOL)."""
    if "sens_pop" not in gold:
        return np.zeros(len(gold["iters"]), dtype=bool)
    return (np.asarray(gold["sens_flag"]) != 0) | (np.asarray(gold["sens_iters"]) > 1) | ~(np.asarray(gold["sens_pop"]) <= SENS_TOL)


# Layers exempt from "flags equal, |d iters| <= 1, populations 1e-6" BY NAME: {case: {layer: reason}}.  A layer that is
# neither listed here nor rounding-sensitive by the fixture's own measurement and misses any of the three fails the test.
# (Layer 152 of the headline slab -- condition number 1e14..1e15, DESIGN.md section 2 -- needs no entry: the fixture's own
# probes mark it rounding-sensitive.)
ALLOW = {
    "ch3oh": {
        108: "the REFERENCE returns NaN populations flagged as converged (its Ng step at iteration 45 solves a numerically "
             "singular 4x4 system and NaN passes `rel_error < min_error`); this build converges to finite populations",
    },
}


def problem(case):
    return CASES[case][0]()


def fixture_path(case):
    return os.path.join(GOLDEN_DIR, f"full_{case}.npz")


def load(case):
    z = np.load(fixture_path(case))
    return {k: z[k] for k in z.files}


def compare(p, pop, st, gold, allow):
    """All-layer comparison of a GPU solve with the reference's committed answer.  Returns (report dict, list of failing
    (layer, what) pairs after the allow-list)."""
    it_g, cv_g = st["iters"].astype(int), st["converged"].astype(int)
    it_r, cv_r, pop_r = gold["iters"].astype(int), gold["converged"].astype(int), gold["pop"]
    L = p.L
    rel = np.full(L, np.nan)
    ref_finite = np.isfinite(pop_r).all(axis=1)
    for l in range(L):
        big = np.isfinite(pop_r[l]) & (pop_r[l] > 1e-30)
        if ref_finite[l] and big.any() and np.isfinite(pop[l]).all():
            rel[l] = np.max(np.abs(pop[l] - pop_r[l])[big] / pop_r[l][big])
    sens = sensitive(gold)
    fails = []
    for l in range(L):
        if l in allow:
            continue
        if sens[l]:
            # The reference's own answer for this layer moves beyond the tolerance when only the rounding of its dense
            # solve changes: asserted at the level such a layer supports -- a finite, normalised answer that agrees with
            # the reference on every level that carries population, and the same outcome unless the reference's own
            # outcome flips under the probes.
            carry = np.isfinite(pop_r[l]) & (pop_r[l] > 1e-8)
            if not np.isfinite(pop[l]).all():
                fails.append((l, "rounding-sensitive layer: output not finite"))
            elif ref_finite[l] and carry.any() and np.max(np.abs(pop[l] - pop_r[l])[carry] / pop_r[l][carry]) > 1e-2:
                fails.append((l, "rounding-sensitive layer: populated levels differ by more than 1e-2"))
            elif cv_g[l] != cv_r[l] and not gold["sens_flag"][l] and not (gold["sens_iters"][l] > 1):
                fails.append((l, f"rounding-sensitive layer: flag {cv_g[l]} vs {cv_r[l]} although the reference's own flag and count are stable"))
            continue
        if cv_g[l] != cv_r[l]:
            fails.append((l, f"flag {cv_g[l]} vs reference {cv_r[l]} (iters {it_g[l]} vs {it_r[l]})"))
            continue
        if cv_r[l] == 1:
            if not ref_finite[l]:
                fails.append((l, "reference output not finite and layer not allow-listed"))
            elif abs(it_g[l] - it_r[l]) > 1:
                fails.append((l, f"iters {it_g[l]} vs {it_r[l]}"))
            elif not (rel[l] <= 1e-6):
                fails.append((l, f"populations differ by {rel[l]:.3e}"))
        elif it_g[l] != it_r[l]:
            fails.append((l, f"non-converged layer stopped at {it_g[l]} vs {it_r[l]} (iteration cap)"))
    both = (cv_g == 1) & (cv_r == 1) & np.isfinite(rel)
    dit = np.abs(it_g - it_r)
    report = dict(
        N=int(p.N), L=int(L), flags_equal=int((cv_g == cv_r).sum()), reference_converged=int(cv_r.sum()),
        gpu_converged=int(cv_g.sum()), iter_diff_le1=int((dit <= 1).sum()), iter_diff_max=int(dit.max()),
        layers_iter_diff_gt1=[(int(l), int(it_g[l]), int(it_r[l])) for l in np.flatnonzero(dit > 1)],
        both_converged=int(both.sum()), pop_rel_err_le_1e6=int((rel[both] <= 1e-6).sum()),
        pop_rel_err_max=float(np.nanmax(rel[both])) if both.any() else None,
        pop_rel_err_median=float(np.nanmedian(rel[both])) if both.any() else None,
        layers_pop_err_gt_1e6=[(int(l), float(rel[l])) for l in np.flatnonzero(both & (rel > 1e-6))],
        layers_reference_not_finite=[int(l) for l in np.flatnonzero(~ref_finite)],
        iters_mean_reference=float(it_r.mean()), iters_max_reference=int(it_r.max()),
        rounding_sensitive_layers=[int(l) for l in np.flatnonzero(sens)],
        rounding_sensitive_detail={int(l): dict(sens_pop=float(gold["sens_pop"][l]), sens_iters=int(gold["sens_iters"][l]),
                                                sens_flag=int(gold["sens_flag"][l]), gpu_pop_err=float(rel[l]),
                                                gpu_iters=int(it_g[l]), ref_iters=int(it_r[l]))
                                   for l in np.flatnonzero(sens)} if "sens_pop" in gold else {},
        outside_parity_all=[(int(l), int(it_g[l]), int(it_r[l]), float(rel[l])) for l in range(L)
                            if cv_g[l] != cv_r[l] or (cv_r[l] == 1 and (abs(it_g[l] - it_r[l]) > 1 or not (rel[l] <= 1e-6)))],
        allow_list={int(k): v for k, v in allow.items()}, failures=[(int(l), w) for l, w in fails])
    return report, fails
