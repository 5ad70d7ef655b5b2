"""GPU (B200): EVERY layer of every BASELINE.json slab at full size -- CUDA path through the C ABI against the committed
answer of the unmodified reference (tests/golden/full_<case>.npz, generated from oracle/_ref by
tests/golden/make_full_slab.py).  Asserted on every layer: identical convergence flag; where the reference converged with
finite output, iteration count within +-1 and populations within 1e-6 relative on levels above 1e-30 (north_star
tolerances); where it hit the iteration cap, the same iteration count.  Exceptions are the layers named, with a reason, in
tests/full_slab_cases.ALLOW -- a new outlier fails the test.  A JSON report per case goes to gpurun_out/ (and from there
to profiles/) when that directory exists."""
import json
import os

import numpy as np
import pytest

import full_slab_cases as F
import golden_cases as G

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _report(case, rep):
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, f"parity_full_slab_{case}.json"), "w") as f:
            json.dump(rep, f, indent=1)


@pytest.mark.parametrize("case", list(F.CASES))
def test_full_slab_every_layer(case):
    import torch
    assert torch.cuda.is_available(), "the gpu-marked tests need a CUDA device"
    from radiative_transfer_b200 import abi, solver
    p = F.problem(case)
    gold = F.load(case)
    assert G.problem_digest(p) == bytes(gold["digest"]).decode(), "fixture was generated for other inputs: re-run tests/golden/make_full_slab.py"
    m = solver.Model(p)
    if F.CASES[case][1] == "chained":
        pop, st = m.calc_molecular_populations(start_mode=abi.RT_START_CHAINED)
    else:
        pop, st = m.calculate_populations()
    m.close()
    assert pop.shape == gold["pop"].shape
    assert np.allclose(pop.sum(axis=1), 1.0, atol=1e-9)
    allow = F.ALLOW.get(case, {})
    rep, fails = F.compare(p, pop, st, gold, allow)
    rep.update(case=case, mode=F.CASES[case][1], reference_cpu_seconds=float(gold["cpu_seconds"]), reference_threads=int(gold["threads"]))
    _report(case, rep)
    assert not fails, f"{case}: {len(fails)} layer(s) outside parity and not allow-listed: {fails[:10]}"
    # the allow-list must not rot: an allow-listed layer that is in full parity again should be taken off the list
    assert len(allow) <= max(2, p.L // 50), "allow-list too long to be an exception list"


@pytest.mark.parametrize("case", ["ch3oh", "h2o", "oh_overlap"])
def test_full_slab_maser_postpass_and_luminosity(case):
    """The post-solve quantities at BASELINE size: maser gains / optical depths
    (transition_data.cpp:210-378) and pump / loss rates + limiting luminosity (maser_luminosity.cpp:7-106) of the six most
    inverted lines, evaluated on the REFERENCE's converged populations of the whole slab (the committed fixture) by the CUDA
    path and by the checker run live on the host (the compiled reference where its .so travelled, else the C restatement):
    the same input on both sides, so the tolerances are those of the small golden cases (1e-10 of the array scale, 1e-9)."""
    import torch
    assert torch.cuda.is_available(), "the gpu-marked tests need a CUDA device"
    from oracle.loader import Checker, ref_available
    from radiative_transfer_b200 import solver
    p = F.problem(case)
    gold = F.load(case)
    pop = np.array(gold["pop"])
    bad = np.flatnonzero(~np.isfinite(pop).all(axis=1))
    for l in bad:  # a layer the reference returned as NaN takes its neighbour's populations: any finite input will do
        pop[l] = pop[l - 1 if l > 0 else l + 1]
    up, low = G.maser_lines(p, pop)
    chk = Checker("ref" if ref_available() else "port")
    want = chk.maser_postpass(p, pop, up, low)
    m = solver.Model(p)
    got = m.maser_postpass(pop, up, low)
    lim = m.lim_luminosity(pop, up, low, beaming_factor=1.0, lay_nb_hg=got["scalars"][:, 4])
    m.close()
    assert (want["inv"] > 0).any(), "the slab must hold inverted lines"
    for k in ("inv", "gain", "scalars", "tau_vs_aspect", "tau_vs_freq"):
        scale = np.max(np.abs(want[k])) + 1e-300
        assert np.max(np.abs(got[k] - want[k])) / scale < 1e-10, k
    want_l = chk.lim_luminosity(p, pop, up, low, beaming_factor=1.0, lay_nb_hg=want["scalars"][:, 4])
    for k in ("lum_arr", "pump_eff_arr", "loss_rate_arr", "pump_rate_arr", "emiss_coeff_arr", "exc_temp_arr", "lum", "tau_sat", "temp_sat"):
        assert np.allclose(lim[k], want_l[k], rtol=1e-9, atol=0.0, equal_nan=True), k


@pytest.mark.parametrize("case,stride", [("oh", 1), ("oh_overlap", 4), ("h2o150", 2), ("h2o", 5), ("ch3oh", 8)])
def test_full_slab_one_step_from_reference_populations(case, stride):
    """One calc_new_pop (iteration_lvg.cpp:87-110: intensities, assembly, dense solve) at BASELINE size from the REFERENCE's
    converged populations, every `stride`-th layer, CUDA path against the checker on the same input: new populations within
    1e-6 (levels above 1e-30) and eq_error within 1e-6 relative (+1e-11, its rounding floor) on every layer that is not rounding-sensitive by the fixture's
    own measurement -- the one-step counterpart of the converged-solve test, free of the iteration's amplification."""
    import torch
    assert torch.cuda.is_available(), "the gpu-marked tests need a CUDA device"
    from oracle.loader import Checker, ref_available
    from radiative_transfer_b200 import solver
    p = F.problem(case)
    gold = F.load(case)
    sens = F.sensitive(gold)
    idx = [l for l in range(0, p.L, stride) if np.isfinite(gold["pop"][l]).all() and not sens[l] and l not in F.ALLOW.get(case, {})]
    assert len(idx) >= 20
    q = p.take(idx)
    pop = np.ascontiguousarray(gold["pop"][idx])
    chk = Checker("ref" if ref_available() else "port")
    want, eq_w = chk.step(q, pop)
    m = solver.Model(q)
    got, eq_g = m.calc_new_pop(pop)
    bnd = m.boundary_layer_populations()
    m.close()
    assert np.allclose(got.sum(axis=1), 1.0, atol=1e-9)
    # boundary_layer_populations (iteration_control.cpp:52-91) of the same layers: the start vectors of the measured mode
    bnd_w = chk.boundary(q)
    for k, l in enumerate(idx):
        big = bnd_w[k] > 1e-30
        assert float(np.max(np.abs(bnd[k] - bnd_w[k])[big] / bnd_w[k][big])) < 1e-7, (case, l, "boundary")
    for k, l in enumerate(idx):
        big = want[k] > 1e-30
        err = float(np.max(np.abs(got[k] - want[k])[big] / want[k][big]))
        assert err < 1e-6, (case, l, err)
        # eq_error = max |e0 - M pop| is a residual AT convergence: rounding of row sums of O(1e3) rates sets its floor
        assert abs(eq_g[k] - eq_w[k]) <= 1e-6 * eq_w[k] + 1e-11, (case, l, eq_g[k], eq_w[k])
