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
