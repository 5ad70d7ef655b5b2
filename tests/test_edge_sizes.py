"""The level counts of the `edge*` parity cases sit on the seams of the solver's size classes (fused_kernels.cu:
fused_lu_variant -- row slots per lane of the register-resident LU: n <= 32, 62, 96, 128, 150; 152..158 LU in shared memory;
above that the blocked multi-kernel LU).  The GPU tests solve them; this pins the level counts the names promise, and that
the CPU checker converges on the smallest ones, so that the seams stay covered when the generator changes."""
import numpy as np
import pytest

from radiative_transfer_b200 import synthetic

SIZES = {"edge4": 4, "edge9": 9, "edge32": 32, "edge33": 33, "edge62": 62, "edge64": 64, "edge98": 98, "edge128": 128,
         "edge136": 136, "edge152": 152, "edge161": 161, "h2o_odd": 131, "h2o_small": 150, "ch3oh_small": 96}


@pytest.mark.parametrize("name,N", sorted(SIZES.items()))
def test_level_counts(name, N):
    p = synthetic.small(name, 2)
    assert p.N == N
    assert p.L == 2


def test_every_size_class_has_a_case():
    def size_class(n):  # mirrors fused_lu_variant + the two paths above it
        n += n & 1
        for k, top in enumerate((32, 62, 96, 128, 150)):
            if n <= top:
                return k + 1
        return 6 if n <= 158 else 7
    assert {size_class(N) for N in SIZES.values()} == {1, 2, 3, 4, 5, 6, 7}


@pytest.mark.parametrize("name", ["edge4", "edge9"])
def test_smallest_cases_converge_on_the_checker(best_checker, name):
    p = synthetic.small(name, 3)
    pop, st, _ = best_checker.solve(p, nthreads=2)
    assert np.all(st["converged"] == 1)
    assert np.allclose(pop.sum(axis=1), 1.0, atol=1e-12)
