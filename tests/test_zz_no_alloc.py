"""GPU (B200), collected LAST: rt_intensity_noalloc -- the allocation-free variant of rt_intensity that keeps its device
scratch in a fixed file-scope device array (radiative_transfer_b200/csrc/noalloc_api.cu; no counterpart in the reference)
-- against the existing entry point rt_intensity, through the C ABI, on calls
in a row that share the array: same kernel, same inputs, so the results must be identical (asserted to 1e-11, the
tolerance of the intensity golden test).  Self-contained: passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _lines(p, n=6):
    iu, il = np.nonzero(np.tril(p.einst, -1) > 0)
    sel = np.linspace(0, iu.shape[0] - 1, min(n, iu.shape[0])).astype(int)
    return [(int(iu[k]), int(il[k])) for k in sel]


def test_noalloc_intensity_matches_rt_intensity_on_consecutive_calls():
    import torch
    assert torch.cuda.is_available(), "the gpu-marked tests need a CUDA device"
    from radiative_transfer_b200 import solver, synthetic
    # a larger molecule first, then a smaller one: the shared array still holds the first call's populations beyond the
    # second molecule's N, which must not matter; the overlap scheme exercises the pair branch and the "not a line" branch
    for name, L in (("ch3oh_small", 3), ("oh_overlap", 3), ("h2o_small", 2)):
        p = synthetic.small(name, L=L)
        m = solver.Model(p)
        b = m.boundary_layer_populations()
        for rep in range(2):                       # two passes in a row over the same handle
            for lay in (0, L - 1):
                for (u, l) in _lines(p):
                    want = m.intensity_calc(lay, u, l, b[lay])
                    got = m.intensity_calc_noalloc(lay, u, l, b[lay])
                    again = m.intensity_calc_noalloc(lay, u, l, b[lay])   # second call in a row, same inputs
                    assert np.isfinite(want)
                    assert abs(got - want) <= 1e-11 * max(abs(want), 1e-300), (name, lay, u, l, got, want)
                    assert again == got
        if name == "oh_overlap":  # a pair that is not a line of the scheme (inside a hyperfine doublet): 0, as rt_intensity
            assert m.intensity_calc_noalloc(0, 1, 0, b[0]) == m.intensity_calc(0, 1, 0, b[0]) == 0.0
        with pytest.raises(solver.RtError):
            m.intensity_calc_noalloc(0, p.N, 0, b[0])  # level index out of range
        m.close()


def test_noalloc_limits_are_declared():
    """The header states the limit; the array is modest (every GPU that loads the library holds it)."""
    import os, re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    txt = open(os.path.join(root, "include", "rt_b200.h")).read()
    mx = int(re.search(r"#define RT_NOALLOC_MAX_LEVELS (\d+)", txt).group(1))
    assert 768 <= mx <= 4096
    from radiative_transfer_b200 import solver
    assert hasattr(solver.lib(), "rt_intensity_noalloc")
