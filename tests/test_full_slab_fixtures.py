"""CPU: the committed full-slab fixtures (tests/golden/full_<case>.npz, answers of the unmodified reference) are valid for
the current generator (input digest), self-consistent, and reproduced by the C restatement (oracle/rt_oracle.c) on a
bounded sample of layers; the comparison logic of the GPU test flags what it must flag."""
import copy

import numpy as np
import pytest

import full_slab_cases as F
import golden_cases as G
from conftest import rel_err


@pytest.mark.parametrize("case", list(F.CASES))
def test_fixture_matches_generator_and_is_sane(case):
    p = F.problem(case)
    gold = F.load(case)
    assert G.problem_digest(p) == bytes(gold["digest"]).decode()
    assert gold["pop"].shape == (p.L, p.N)
    conv = gold["converged"] == 1
    fin = np.isfinite(gold["pop"]).all(axis=1)
    assert conv.mean() >= 0.9, "the slab must measure the solver, not the retry policy"
    assert np.allclose(gold["pop"][conv & fin].sum(axis=1), 1.0, atol=1e-9)
    # layers the reference returns as non-finite must be on the allow-list (they cannot be compared)
    for l in np.flatnonzero(~fin):
        assert int(l) in F.ALLOW.get(case, {}), (case, int(l))
    # the rounding-sensitivity map is there, and the slab is mostly made of layers that ARE held to 1e-6
    assert "sens_pop" in gold and len(gold["sens_probes"]) >= 3
    assert F.sensitive(gold).mean() <= 0.08, "too many rounding-sensitive layers: the slab would test little"
    assert len(F.ALLOW.get(case, {})) <= 2


@pytest.mark.parametrize("case,layers", [("oh", [0, 50, 99]), ("oh_overlap", [0, 399]), ("h2o150", [10, 150]), ("ch3oh", [60])])
def test_restatement_reproduces_fixture_on_sample(port, case, layers):
    p = F.problem(case)
    gold = F.load(case)
    for l in layers:
        q = p.sliced(l, l + 1)
        pop, st, _ = port.solve(q)
        assert st["converged"][0] == gold["converged"][l]
        assert abs(int(st["iters"][0]) - int(gold["iters"][l])) <= 1
        assert rel_err(pop[0], gold["pop"][l], 1e-30) < 1e-6


def test_compare_flags_outliers():
    case = "oh"
    p = F.problem(case)
    gold = F.load(case)
    st = np.zeros(p.L, dtype=[("iters", "<i4"), ("converged", "<i4")])
    st["iters"], st["converged"] = gold["iters"], gold["converged"]
    pop = gold["pop"].copy()
    rep, fails = F.compare(p, pop, st, gold, {})
    assert not fails and rep["flags_equal"] == p.L
    pop[7, 3] *= 1.0 + 5e-6
    st["iters"][11] += 2
    st["converged"][13] ^= 1
    rep, fails = F.compare(p, pop, st, gold, {})
    assert sorted(l for l, _ in fails) == [7, 11, 13]
    rep, fails = F.compare(p, pop, st, gold, {7: "x", 11: "y", 13: "z"})
    assert not fails
